#!/usr/bin/env python3
"""Hot SASS instructions of one kernel in an .ncu-rep: tools/ncu_hot.py rep kernel-regex [top]"""
import csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + kern, '--print-source', 'sass'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
hdr = rows[hi]
body = []
for r in rows[hi + 1:]:
    if len(r) < len(hdr) or r[0] == 'Address' or not r[0].startswith('0x'): break
    body.append(r)
S = hdr.index('# Samples'); I = hdr.index('Instructions Executed')
tot_s = sum(int(r[S]) for r in body); tot_i = sum(int(r[I]) for r in body)
print(f'{len(body)} instrs, samples {tot_s}, warp-inst {tot_i}')
stalls = [c for c in hdr if c.startswith('stall_') and 'Not Issued' not in c]
agg = {c: sum(int(r[hdr.index(c)]) for r in body) for c in stalls}
print('stall mix:', ', '.join(f'{k[6:]} {100*v/max(tot_s,1):.0f}%' for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
idx = sorted(range(len(body)), key=lambda i: -int(body[i][S]))[:top]
for i in sorted(idx):
    r = body[i]
    st = sorted(((int(r[hdr.index(c)]), c[6:]) for c in stalls), reverse=True)[:2]
    print(f'{i:5d} {100*int(r[S])/max(tot_s,1):5.1f}% smp  {int(r[I]):9d} exec  {r[1].strip()[:70]:70s} {st[0][1]}:{st[0][0]} {st[1][1]}:{st[1][0]}')
