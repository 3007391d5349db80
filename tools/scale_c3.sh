#!/bin/bash
# config 3 (population 5000 pairs, strong scaling) at the GPU counts given, e.g. tools/scale_c3.sh tag "4 8"
tag=${1:-sc}
mkdir -p gpurun_out
out=gpurun_out/${tag}_configs.jsonl; : > $out
port=29600
for n in $2; do
  port=$((port+1))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port tools/run_configs.py --config 3 --steps 3 >> $out 2> gpurun_out/${tag}_c3_$n.err; echo "c3 n=$n rc=$?"
done
grep config $out | cut -c1-330
