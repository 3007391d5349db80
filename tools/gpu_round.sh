#!/bin/bash
# one GPU call: parity tests, smoke, default bench (+ exact / fast lines), ncu launch list and a full capture of the member GEMMs
tag=${1:-cur}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 -x > gpurun_out/${tag}_gputests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${tag}_gputests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${tag}_smoke.log
timeout 600 python bench.py > gpurun_out/${tag}_bench_hybrid.json 2> gpurun_out/${tag}_bench_hybrid.err; echo "bench rc=$?"; tail -2 gpurun_out/${tag}_bench_hybrid.err
timeout 600 python bench.py --precision exact --no-extras > gpurun_out/${tag}_bench_exact.json 2> gpurun_out/${tag}_bench_exact.err; tail -2 gpurun_out/${tag}_bench_exact.err
timeout 300 python bench.py --precision fast --no-extras > gpurun_out/${tag}_bench_fast.json 2> gpurun_out/${tag}_bench_fast.err; tail -2 gpurun_out/${tag}_bench_fast.err
if [ "${2:-ncu}" = "ncu" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 1 --warmup 1 --no-extras > gpurun_out/${tag}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_member_gemm|k_gradient|k_refine" -s 60 -c 6 -o gpurun_out/${tag}_full -f python bench.py --steps 1 --warmup 1 --population 296 --no-extras > gpurun_out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_gradient|k_opt_update" -s 3 -c 3 -o gpurun_out/${tag}_full_master -f python bench.py --steps 1 --warmup 1 --population 148 --cpu-seconds 1 > gpurun_out/${tag}_ncu_full_master.log 2>&1; echo "ncu full master rc=$?"
fi
python - <<PY
import json
for m in ("hybrid","exact","fast"):
    try:
        d=json.loads(open('gpurun_out/${tag}_bench_%s.json'%m).read().strip().splitlines()[-1]); r=d['roofline']
        sh=r['kernel_time_share']; ms=d['ms_per_step']
        print(m,'evals/s %.0f ms %.1f e2e %.0f clk %s logit_ms %.2f lstm_ms %.2f frac %.3f shares %s'%(d['value'],ms,d['e2e']['value'],d['clocks']['sm_mhz'],r['avg_launch_ms'],sh['lstm']*ms/17,r['frac'],{k:round(v,3) for k,v in sh.items()}))
    except Exception as e: print(m,'failed',e)
PY
