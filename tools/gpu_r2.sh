#!/bin/bash
# one GPU call of round 2: parity tests, smoke, default bench line (with per-phase sections and the config-3 theta hash written when asked)
tag=${1:-r2}
mode=${2:-all}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv,noheader > gpurun_out/${tag}_gpu.txt 2>&1
if [ "$mode" != bench ]; then
  timeout 1500 python -m pytest tests -m gpu -q --timeout 900 $PYTEST_ARGS > gpurun_out/${tag}_gputests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/${tag}_gputests.log
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${tag}_smoke.log
fi
timeout 900 python bench.py --sections $BENCH_ARGS > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench.err
cp tests/golden/config3_theta_hash.json gpurun_out/${tag}_config3_theta_hash.json 2>/dev/null
python tools/show_bench.py gpurun_out/${tag}_bench.json $tag
