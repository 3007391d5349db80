#!/bin/bash
# timing experiments with the NES_DEBUG bits (results may be WRONG for non-zero values; timing only): usage exp_debug.sh "0 8 0 8"
for d in ${1:-0 8 0 8}; do
  NES_DEBUG=$d timeout 300 python bench.py --no-extras --steps 4 --warmup 3 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; sh=r['kernel_time_share']; ms=d['ms_per_step']
print('DEBUG=$d evals/s %.0f ms %.1f clk %s logit_ms %.2f lstm_ms %.2f mat_ms %.2f fin_ms %.2f'%(d['value'],ms,d['clocks']['sm_mhz'],r['avg_launch_ms'],sh['lstm']*ms/17,sh['materialise']*ms,sh['finalize']*ms))"
done
nvidia-smi --query-gpu=power.limit,power.draw,temperature.gpu,clocks.sm,clocks.mem --format=csv
