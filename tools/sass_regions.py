"""Dynamic instruction counts / stall samples per loop of a kernel, from the `--all-sass` dump of tools/ncu_summary.py.
    python tools/sass_regions.py X_all_sass.csv [min_share_pct]
Loops are found from backward branches; each is listed with its executed warp instructions, its share of the kernel,
its stall samples and its dominant stall reasons and opcodes (nested loops are listed on their own and inside parents)."""
import collections, csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))[1:]
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 2.0
base = int(rows[0][0], 16)
ins = [(int(r[0], 16) - base, float(r[1]), float(r[2]), r[3], r[4]) for r in rows]
tot, tots = sum(i[1] for i in ins), sum(i[2] for i in ins)
print(f'total warp instructions {tot:.0f}, stall samples {tots:.0f}, {len(ins)} SASS instructions')
loops = set()
for a, e, s, st, t in ins:
    if 'BRA' in t:
        m = re.search(r'0x([0-9a-f]+)', t)
        if m and int(m.group(1), 16) - base < a:
            loops.add((int(m.group(1), 16) - base, a))
for lo, hi in sorted(loops):
    sel = [i for i in ins if lo <= i[0] <= hi]
    e, s = sum(i[1] for i in sel), sum(i[2] for i in sel)
    if 100 * e / tot < thr:
        continue
    ops, stl = collections.Counter(), collections.Counter()
    for i in sel:
        m = re.match(r'\s*(?:@!?U?P\d+\s+)?([A-Z0-9_]+)', i[4])
        ops[m.group(1) if m else '?'] += i[1]
        if i[3]:
            stl[i[3].replace('stall_', '')] += i[2]
    print(f'loop {lo:#x}..{hi:#x} ({len(sel)} SASS): {100 * e / tot:.1f}% of instructions, {100 * s / tots:.1f}% of stall samples')
    print('   ops:    ' + ', '.join(f'{k} {100 * v / e:.1f}%' for k, v in ops.most_common(14)))
    print('   stalls (by the top reason of each instruction): ' + ', '.join(f'{k} {100 * v / max(s, 1):.0f}%' for k, v in stl.most_common(6)))
