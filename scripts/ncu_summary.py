#!/usr/bin/env python
"""Side-by-side key metrics of .ncu-rep files: python scripts/ncu_summary.py a.ncu-rep b.ncu-rep ..."""
import csv, io, subprocess, sys
KEYS = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__warps_eligible.avg.per_cycle_active',
        'sm__icc_request_hit_rate.pct',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'dram__bytes_read.sum', 'dram__bytes_write.sum']


def raw(path):
    txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    return dict(zip(rows[0], rows[2] if len(rows) > 2 else rows[1]))


docs = [raw(p) for p in sys.argv[1:]]
print("metric," + ",".join(p.split('/')[-1].replace('.ncu-rep', '') for p in sys.argv[1:]))
for k in KEYS:
    print(k + "," + ",".join(d.get(k, '') for d in docs))
