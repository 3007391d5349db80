#!/bin/bash
# A/B of one environment switch on one box: tools/ab_env.sh VAR A_VALUE B_VALUE [bench args]; prints one summary line per leg (A B A B)
var=$1; a=$2; b=$3; shift 3
mkdir -p gpurun_out
for leg in A B A B; do
  if [ $leg = A ]; then v=$a; else v=$b; fi
  env $var=$v timeout 600 python bench.py --steps 10 --warmup 3 --no-extras --nes-pairs 0 "$@" > gpurun_out/ab_${var}_${leg}.json 2> gpurun_out/ab_${var}_${leg}.err
  python tools/show_bench.py gpurun_out/ab_${var}_${leg}.json "$var=$v"
done
