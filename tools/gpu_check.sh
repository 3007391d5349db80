#!/bin/bash
# one GPU call: parity tests, then the bench in both precisions; summaries into gpurun_out/<tag>_*
tag=${1:-cur}
timeout 600 python -m pytest tests -m gpu -q --timeout 300 -x > gpurun_out/${tag}_gputests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${tag}_gputests.log
timeout 600 python bench.py --no-extras > gpurun_out/${tag}_bench_hybrid.json 2> gpurun_out/${tag}_bench_hybrid.err; tail -2 gpurun_out/${tag}_bench_hybrid.err
timeout 600 python bench.py --precision exact --no-extras > gpurun_out/${tag}_bench_exact.json 2> gpurun_out/${tag}_bench_exact.err; echo "bench rc=$?"; tail -2 gpurun_out/${tag}_bench_exact.err
timeout 300 python bench.py --precision fast --no-extras > gpurun_out/${tag}_bench_fast.json 2> gpurun_out/${tag}_bench_fast.err; tail -2 gpurun_out/${tag}_bench_fast.err
python - <<PY
import json
for m in ("hybrid","exact","fast"):
    try:
        d=json.loads(open('gpurun_out/${tag}_bench_%s.json'%m).read().strip().splitlines()[-1]); r=d['roofline']
        sh=r['kernel_time_share']; ms=d['ms_per_step']
        print(m,'evals/s %.0f ms %.1f e2e %.0f clk %s logit_ms %.2f lstm_ms %.2f frac %.3f shares %s'%(d['value'],ms,d['e2e']['value'],d['clocks']['sm_mhz'],r['avg_launch_ms'],sh['lstm']*ms/17,r['frac'],{k:round(v,3) for k,v in sh.items()}))
    except Exception as e: print(m,'failed',e)
PY
