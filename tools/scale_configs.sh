#!/bin/bash
# on an 8-GPU box: config 3 (NIC-NES, population 5000 pairs, strong scaling) at 1/2/4/8 GPUs, config 4 (LayerNorm shapes, population
# 10000 rollouts, 512 rows) at 8 GPUs, and the default bench line at 8 GPUs
tag=${1:-sc}
mkdir -p gpurun_out
out=gpurun_out/${tag}_configs.jsonl; : > $out
port=29500
for n in 1 2 4 8; do
  port=$((port+1))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port tools/run_configs.py --config 3 --steps 3 >> $out 2> gpurun_out/${tag}_c3_$n.err; echo "c3 n=$n rc=$?"
done
port=$((port+1))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port tools/run_configs.py --config 4 --pairs 5000 --steps 3 >> $out 2> gpurun_out/${tag}_c4_8.err; echo "c4 n=8 rc=$?"
port=$((port+1))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 8 --no-extras > gpurun_out/${tag}_bench_8gpu.json 2> gpurun_out/${tag}_bench_8gpu.err; echo "bench n=8 rc=$?"
cat $out | cut -c1-400
cut -c1-400 gpurun_out/${tag}_bench_8gpu.json
