#!/usr/bin/env python3
"""Executed-instruction mix by opcode of one kernel in an .ncu-rep: tools/ncu_opmix.py rep kernel-regex"""
import csv, io, subprocess, sys, collections
rep, kern = sys.argv[1], sys.argv[2]
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + kern, '--print-source', 'sass'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
hdr = rows[hi]; I = hdr.index('Instructions Executed')
mix = collections.Counter(); tot = 0; seen = set()
for r in rows[hi + 1:]:
    if len(r) < len(hdr) or not r[0].startswith('0x'): break
    if r[0] in seen: break
    seen.add(r[0])
    t = r[1].strip().split()
    op = t[1] if t[0].startswith('@') else t[0]
    op = op.split('.')[0] + ('.' + op.split('.')[1] if op.startswith(('LD', 'ST', 'ATOM', 'VIMNMX')) and '.' in op else '')
    mix[op] += int(r[I]); tot += int(r[I])
print('total warp-inst', tot)
for op, n in mix.most_common(28): print(f'{op:16s} {n:12d} {100*n/tot:5.1f}%')
