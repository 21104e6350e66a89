#!/usr/bin/env python3
"""Per-source-line samples / executed instructions of one kernel: tools/ncu_lines.py rep kernel-regex [min_samples]"""
import csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
mins = int(sys.argv[3]) if len(sys.argv) > 3 else 20
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + kern, '--print-source', 'cuda,sass'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
cur = None; tot = 0; out = []
for r in rows:
    if len(r) == 2 and r[0] == 'File Name': cur = r[1].split('/')[-1]; continue
    if len(r) < 8 or r[0] in ('Line No', ''): continue
    try: ln = int(r[0]); smp = int(r[6]); ex = int(r[7])
    except ValueError: continue
    tot += smp; out.append((cur, ln, smp, ex, r[1].strip()))
print('total samples', tot)
for f, ln, smp, ex, src in out:
    if smp >= mins: print(f'{f}:{ln:4d} {100*smp/max(tot,1):5.1f}% {ex:10d}  {src[:110]}')
