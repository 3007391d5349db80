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
    kk = ' '.join('%s %.2fms %.0f%s(%.2f)' % (n.replace('k_member_', ''), v.get('avg_launch_ms', v.get('ms_per_wave', 0)), v['achieved'], v['unit'], v['frac']) for n, v in ks.items())
    print('%s evals/s %.0f ms %.1f e2e %.0f clk %s | %s | shares %s | hyb %s' % (tag, d['value'], ms, d['e2e']['value'], d['clocks']['sm_mhz'], kk,
          {k: round(v * ms, 1) for k, v in sh.items()}, d.get('hybrid_stats', {}).get('rows_overflowed')))
    if 'nes_config3' in d:
        n3 = d['nes_config3']
        print('   config3: %d GPUs %.1f ms / iteration, %.0f rollouts/s, theta_agree %s (ranks %s, golden %s)' % (n3['n_gpus'], n3['iteration_ms'], n3['rollouts_per_s'], n3['theta_agree'], n3['theta_identical_on_all_ranks'], n3['theta_equals_1rank_golden']))
    if 'sections_ms_per_rank' in d:
        for r in d['sections_ms_per_rank']:
            print('   ', r)
except Exception as e:
    print(tag, 'failed', repr(e))
