"""Opcode histogram of one kernel from `cuobjdump -sass` output (static counts; the hot loop is `#pragma unroll 1`, so
the static mix of the rollout body is what executes per step).  usage: sass_stats.py file.o 'mangled-substring' [...]"""
import collections
import re
import subprocess
import sys


def functions(path):
    out = subprocess.run(['cuobjdump', '-sass', path], capture_output=True, text=True, check=True).stdout
    cur, funcs = None, {}
    for line in out.splitlines():
        m = re.search(r'Function : (\S+)', line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
            continue
        m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)', line)
        if m and cur:
            funcs[cur].append(m.group(3))
    return funcs


if __name__ == '__main__':
    fs = functions(sys.argv[1])
    for pat in sys.argv[2:]:
        for name, ops in fs.items():
            if pat in name:
                c = collections.Counter(o.split('.')[0] for o in ops)
                tot = sum(c.values())
                print(f'{name}: {tot} instructions')
                print('  ' + ', '.join(f'{k} {v}' for k, v in c.most_common(24)))
