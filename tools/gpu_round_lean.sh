#!/bin/bash
# one GPU call, small outputs: parity tests, smoke, default bench line, ncu launch list, --set full captures summarised ON THE BOX
# (tools/ncu_full.py) so that only text comes back; the dense-kernel report itself is kept
tag=${1:-cur}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 -x > gpurun_out/${tag}_gputests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${tag}_gputests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${tag}_smoke.log
timeout 600 python bench.py > gpurun_out/${tag}_bench_hybrid.json 2> gpurun_out/${tag}_bench_hybrid.err; echo "bench rc=$?"; tail -2 gpurun_out/${tag}_bench_hybrid.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 1 --warmup 1 --no-extras > gpurun_out/${tag}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_member_dense" -s 10 -c 4 -o gpurun_out/${tag}_full_dense -f python bench.py --steps 1 --warmup 1 --no-extras > gpurun_out/${tag}_ncu_full_dense.log 2>&1; echo "ncu full dense rc=$?"
timeout 900 ncu --set full --clock-control none -k regex:"k_materialise|k_refine|k_finalize|k_member_gemm|k_row_margin" -c 14 -o gpurun_out/${tag}_full_aux -f python bench.py --steps 1 --warmup 1 --no-extras > gpurun_out/${tag}_ncu_full_aux.log 2>&1; echo "ncu full aux rc=$?"
timeout 900 ncu --set full --clock-control none -k regex:"k_evolve_batch|k_rowmax_theta_batch" -c 2 -o gpurun_out/${tag}_full_evolve -f python bench.py --steps 1 --warmup 1 --no-extras > gpurun_out/${tag}_ncu_full_evolve.log 2>&1; echo "ncu full evolve rc=$?"
timeout 600 ncu --set full --clock-control none -k regex:"k_gradient|k_opt_update" -s 3 -c 3 -o gpurun_out/${tag}_full_master -f python bench.py --steps 1 --warmup 1 --population 148 --cpu-seconds 1 > gpurun_out/${tag}_ncu_full_master.log 2>&1; echo "ncu full master rc=$?"
for k in dense aux evolve master; do
  python tools/ncu_full.py gpurun_out/${tag}_full_$k.ncu-rep > gpurun_out/${tag}_full_${k}_summary.txt 2>&1
  [ "$k" != dense ] && rm -f gpurun_out/${tag}_full_$k.ncu-rep
done
python tools/show_bench.py gpurun_out/${tag}_bench_hybrid.json hybrid
du -sh gpurun_out
