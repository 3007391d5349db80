#!/bin/bash
# tests + bench + ncu launch list + ncu full capture of the dense member GEMMs at the bench population
tag=${1:-p}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 -x > gpurun_out/${tag}_gputests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${tag}_gputests.log
timeout 600 python bench.py --no-extras > gpurun_out/${tag}_bench_hybrid.json 2> gpurun_out/${tag}_bench_hybrid.err; echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench_hybrid.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 1 --warmup 1 --no-extras > gpurun_out/${tag}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_member_dense|k_materialise\b|k_refine|k_finalize" -s 45 -c 8 -o gpurun_out/${tag}_full -f python bench.py --steps 1 --warmup 1 --no-extras > gpurun_out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${tag}_bench_hybrid.json').read().strip().splitlines()[-1]); r=d['roofline']
    sh=r['kernel_time_share']; ms=d['ms_per_step']
    print('hybrid evals/s %.0f ms %.1f e2e %.0f clk %s logit_ms %.2f lstm_ms %.2f frac %.3f shares %s hyb %s'%(d['value'],ms,d['e2e']['value'],d['clocks']['sm_mhz'],r['avg_launch_ms'],sh['lstm']*ms/17,r['frac'],{k:round(v,3) for k,v in sh.items()}, d['hybrid_stats']))
except Exception as e: print('failed',e)
PY
