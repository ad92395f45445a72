"""Per-CUDA-source-line instruction counts / stall samples from an `ncu --set full --import-source on` report.
    python tools/ncu_lines.py report.ncu-rep [top] [out.csv]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out[out.find('"'):])))
cur, hdr, agg = None, None, []
for r in rows:
    if not r: continue
    if r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r[0] == 'Function Name': continue
    if r[0] == 'Line No': hdr = r; ci = r.index('Instructions Executed'); cs = r.index('# Samples'); continue
    if hdr and len(r) > ci and r[2] == '-' and r[0].isdigit():
        try: agg.append((float(r[ci]), float(r[cs]), cur, int(r[0]), r[1].strip()))
        except ValueError: pass
tot = sum(a[0] for a in agg); tots = sum(a[1] for a in agg)
agg.sort(key=lambda a: -a[0])
lines = [('inst_executed', 'pct', 'stall_samples', 'pct', 'file', 'line', 'source')]
for a in agg[:top]:
    lines.append((int(a[0]), round(100 * a[0] / tot, 2), int(a[1]), round(100 * a[1] / max(tots, 1), 2), a[2], a[3], a[4][:110]))
if len(sys.argv) > 3:
    with open(sys.argv[3], 'w', newline='') as f: csv.writer(f).writerows(lines)
for l in lines: print(*l, sep=' | ')
print('total warp inst', tot, 'samples', tots)
