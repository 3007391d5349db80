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
        'smsp__inst_executed.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active']
idx = {h: i for i, h in enumerate(hdr)}
for r in rows[2:]:
    print('----')
    for w in want:
        if w in idx:
            print('%-70s %s %s' % (w, r[idx[w]], units[idx[w]]))
