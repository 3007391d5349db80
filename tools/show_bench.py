#!/usr/bin/env python
"""one-line summary of a bench.py JSON line (file argument or stdin)"""
import json
import sys
txt = open(sys.argv[1]).read() if len(sys.argv) > 1 and sys.argv[1] != '-' else sys.stdin.read()
tag = sys.argv[2] if len(sys.argv) > 2 else ''
try:
    d = json.loads(txt.strip().splitlines()[-1])
    r = d['roofline']; sh = r['kernel_time_share']; ms = d['ms_per_step']
    ks = {r['kernel'].split(' ')[0]: r}
    ks.update({v['kernel'].split(' ')[0]: v for v in r.get('other_kernels', {}).values()})
    kk = ' '.join('%s %.2fms %.0f%s(%.2f)' % (n.replace('k_member_', ''), v['avg_launch_ms'], v['achieved'], v['unit'], v['frac']) for n, v in ks.items())
    print('%s evals/s %.0f ms %.1f e2e %.0f clk %s | %s | shares %s | hyb %s' % (tag, d['value'], ms, d['e2e']['value'], d['clocks']['sm_mhz'], kk,
          {k: round(v * ms, 1) for k, v in sh.items()}, d.get('hybrid_stats', {}).get('rows_overflowed')))
except Exception as e:
    print(tag, 'failed', repr(e))
