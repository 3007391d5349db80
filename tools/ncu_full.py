#!/usr/bin/env python
"""print the headline metrics of every kernel in an `ncu --set full` report (reads `ncu -i X --page raw --csv`)"""
import csv
import subprocess
import sys
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__bytes_read.sum.per_second',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_tensor.sum',
        'l1tex__m_xbar2l1tex_read_bytes.sum', 'l1tex__m_xbar2l1tex_read_bytes.sum.per_second', 'lts__t_sector_hit_rate.pct',
        'sm__cycles_elapsed.max', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'launch__grid_size', 'launch__registers_per_thread',
        'smsp__inst_executed.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__warps_eligible.avg.per_cycle_active']
idx = {h: i for i, h in enumerate(hdr)}
# warp-state sampling (why warps do not issue): every smsp__average_warp*_issue_stalled_*_per_issue_active ratio the report holds, largest first
stall = [h for h in hdr if 'issue_stalled' in h and h.endswith('per_issue_active.ratio') and 'not_issued' not in h]
for r in rows[2:]:
    print('----')
    for w in want:
        if w in idx:
            print('%-70s %s %s' % (w, r[idx[w]], units[idx[w]]))
    top = sorted(((float(r[idx[h]].replace(',', '') or 0), h) for h in stall), reverse=True)[:5]
    for v, h in top:
        print('%-70s %.2f' % (h.replace('smsp__average_', '').replace('_per_issue_active.ratio', ''), v))
