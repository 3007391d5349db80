#!/bin/bash
# one GPU call: parity tests, smoke, default bench (+ exact / fast lines, reference arm), ncu launch list and full captures
tag=${1:-cur}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 -x > gpurun_out/${tag}_gputests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${tag}_gputests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${tag}_smoke.log
timeout 600 python bench.py > gpurun_out/${tag}_bench_hybrid.json 2> gpurun_out/${tag}_bench_hybrid.err; echo "bench rc=$?"; tail -2 gpurun_out/${tag}_bench_hybrid.err
timeout 600 python bench.py --precision exact --no-extras > gpurun_out/${tag}_bench_exact.json 2> gpurun_out/${tag}_bench_exact.err; tail -2 gpurun_out/${tag}_bench_exact.err
timeout 300 python bench.py --precision fast --no-extras > gpurun_out/${tag}_bench_fast.json 2> gpurun_out/${tag}_bench_fast.err; tail -2 gpurun_out/${tag}_bench_fast.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2> gpurun_out/${tag}_bench_reference.err; echo "reference rc=$?"; cut -c1-300 gpurun_out/${tag}_bench_reference.json
if [ "${2:-ncu}" = "ncu" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 1 --warmup 1 --no-extras > gpurun_out/${tag}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_member_dense" -s 10 -c 4 -o gpurun_out/${tag}_full_dense -f python bench.py --steps 1 --warmup 1 --no-extras > gpurun_out/${tag}_ncu_full_dense.log 2>&1; echo "ncu full dense rc=$?"
timeout 900 ncu --set full --clock-control none -k regex:"k_materialise|k_refine|k_finalize|k_member_gemm|k_row_margin" -c 16 -o gpurun_out/${tag}_full_aux -f python bench.py --steps 1 --warmup 1 --no-extras > gpurun_out/${tag}_ncu_full_aux.log 2>&1; echo "ncu full aux rc=$?"
timeout 900 ncu --set full --clock-control none -k regex:"k_evolve_batch|k_rowmax_theta_batch" -c 2 -o gpurun_out/${tag}_full_evolve -f python bench.py --steps 1 --warmup 1 --no-extras > gpurun_out/${tag}_ncu_full_evolve.log 2>&1; echo "ncu full evolve rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_gradient|k_opt_update" -s 3 -c 3 -o gpurun_out/${tag}_full_master -f python bench.py --steps 1 --warmup 1 --population 148 --cpu-seconds 1 > gpurun_out/${tag}_ncu_full_master.log 2>&1; echo "ncu full master rc=$?"
fi
for m in hybrid exact fast; do python tools/show_bench.py gpurun_out/${tag}_bench_$m.json $m; done
