#!/bin/bash
# bench sweep over the producer tuning knobs; prints one line per setting
for cfg in "0 0" "0 9" "0 11" "0 0"; do
  set -- $cfg
  for prec in exact fast; do
    NES_PF=$1 NES_HINTS=$2 timeout 300 python bench.py --no-extras --precision $prec --steps 3 --warmup 3 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; sh=r['kernel_time_share']; ms=d['ms_per_step']
print('PF=$1 HINTS=$2 $prec evals/s %.0f ms %.1f clk %s logit_ms %.2f lstm_ms %.2f'%(d['value'],ms,d['clocks']['sm_mhz'],r['avg_launch_ms'],sh['lstm']*ms/17))"
  done
done
