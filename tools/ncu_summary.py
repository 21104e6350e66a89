#!/usr/bin/env python3
"""Condense an .ncu-rep (ncu --set full) into a small CSV for profiles/: one row per profiled launch.
usage: tools/ncu_summary.py gpurun_out/prof.ncu-rep profiles/name.csv"""
import csv, io, subprocess, sys
WANT = ['ID', 'Kernel Name', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'lts__t_sectors.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'launch__waves_per_multiprocessor',
        'smsp__average_warp_latency_issue_stalled_short_scoreboard.ratio', 'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio']
raw = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
cols = [(w, hdr.index(w)) for w in WANT if w in hdr]
with open(sys.argv[2], 'w', newline='') as f:
    w = csv.writer(f)
    w.writerow([c for c, _ in cols]); w.writerow([units[i] for _, i in cols])
    for r in rows[2:]:
        w.writerow([r[i].split('(')[0] if c == 'Kernel Name' else r[i] for c, i in cols])
for r in rows[2:]:
    d = {c: r[i] for c, i in cols}
    print(f"{d['Kernel Name'].split('(')[0]:18s} {float(d['gpu__time_duration.sum']):9.1f} {units[hdr.index('gpu__time_duration.sum')]}  "
          f"dram {d['dram__bytes_read.sum']}+{d['dram__bytes_write.sum']} {units[hdr.index('dram__bytes_read.sum')]}  "
          f"issue {float(d['smsp__issue_active.avg.pct_of_peak_sustained_active']):.0f}%  inst {float(d['smsp__inst_executed.sum'])/1e6:.1f}M")
