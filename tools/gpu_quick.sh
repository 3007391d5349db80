#!/bin/bash
# quick GPU loop: parity tests, then the hybrid bench line
tag=${1:-q}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/${tag}_gputests.log 2>&1; echo "tests rc=$?"; tail -15 gpurun_out/${tag}_gputests.log
timeout 600 python bench.py --no-extras > gpurun_out/${tag}_bench_hybrid.json 2> gpurun_out/${tag}_bench_hybrid.err; echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench_hybrid.err
python tools/show_bench.py gpurun_out/${tag}_bench_hybrid.json hybrid
