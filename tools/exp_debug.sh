#!/bin/bash
# timing experiments with the NES_DEBUG bits (results may be WRONG for non-zero values; timing only): usage exp_debug.sh "0 8 0 8"
for d in ${1:-0 8 0 8}; do
  NES_DEBUG=$d timeout 300 python bench.py --no-extras --steps 4 --warmup 3 2>/dev/null | python tools/show_bench.py - DEBUG=$d
done
nvidia-smi --query-gpu=power.limit,power.draw,temperature.gpu,clocks.sm,clocks.mem --format=csv
