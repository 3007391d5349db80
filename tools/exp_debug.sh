#!/bin/bash
# timing experiments with the NES_DEBUG bits of k_member_gemm (results are WRONG for non-zero values; timing only)
for d in 0 2 4 6 1; do
  NES_DEBUG=$d timeout 300 python bench.py --no-extras --steps 3 --warmup 2 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; sh=r['kernel_time_share']; ms=d['ms_per_step']
print('DEBUG=$d evals/s %.0f ms %.1f clk %s logit_ms %.2f lstm_ms %.2f'%(d['value'],ms,d['clocks']['sm_mhz'],r['avg_launch_ms'],sh['lstm']*ms/17))"
done
