#!/bin/bash
# tests + bench + ncu launch list + ncu full capture of the dense member GEMMs at the bench population
tag=${1:-p}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 -x > gpurun_out/${tag}_gputests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${tag}_gputests.log
timeout 600 python bench.py --no-extras > gpurun_out/${tag}_bench_hybrid.json 2> gpurun_out/${tag}_bench_hybrid.err; echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench_hybrid.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 1 --warmup 1 --no-extras > gpurun_out/${tag}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_member_dense|k_materialise\b|k_refine|k_finalize" -s 45 -c 8 -o gpurun_out/${tag}_full -f python bench.py --steps 1 --warmup 1 --no-extras > gpurun_out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
python tools/show_bench.py gpurun_out/${tag}_bench_hybrid.json hybrid
