#!/bin/bash
# one 8-GPU call: headline bench with per-rank sections (weak scaling) + BASELINE configs[2] inside it, configs[3] (bu_xent shapes, LN, 512 rows)
# at population 10 000 pairs, configs[4] = the K7 / K8 sweep with the parameter axis sharded over the ranks
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port"
$TR 29521 bench.py --gpus 8 --steps 5 --warmup 3 --sections > gpurun_out/c1_bench_8gpu.json 2> gpurun_out/c1_bench_8gpu.err; echo "bench rc=$?"
python tools/show_bench.py gpurun_out/c1_bench_8gpu.json n8
$TR 29522 tools/run_configs.py --config 4 --pairs 10000 --steps 2 2> gpurun_out/c1_cfg4.err | tail -1 | tee gpurun_out/c1_cfg4_8gpu.jsonl
$TR 29523 tools/run_configs.py --config 5 2> gpurun_out/c1_cfg5.err | grep config > gpurun_out/c1_cfg5_8gpu.jsonl; echo "sweep rc=$? lines $(wc -l < gpurun_out/c1_cfg5_8gpu.jsonl)"; tail -3 gpurun_out/c1_cfg5_8gpu.jsonl
