#!/bin/bash
# profiling call of round 2 (one GPU): the plain bench command first (must exit 0), then the ncu launch list of the same command and
# `--set full` captures of the top kernels, summarised ON THE BOX (tools/ncu_full.py / ncu_agg.py) so that only text comes back
tag=${1:-r2}
CMD="python bench.py --steps 1 --warmup 1 --no-extras --nes-pairs 0"
mkdir -p gpurun_out
timeout 600 $CMD > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err; rc=$?; echo "plain rc=$rc"; [ $rc -ne 0 ] && exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
python tools/ncu_agg.py gpurun_out/${tag}_launches.csv > gpurun_out/${tag}_launches_summary.txt 2>&1; head -14 gpurun_out/${tag}_launches_summary.txt
# generation 0 of the bench is the fresh-model one (scaled kernels) and every generation first scores its 2 elite candidates with tiny
# launches: the skip counts (-s) land on the full-size launches of generation 1 (indices read off the launch list)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_lstm_pair|k_member_dense" -s 104 -c 4 -o gpurun_out/${tag}_full_dense -f $CMD > gpurun_out/${tag}_ncu_full_dense.log 2>&1; echo "ncu full dense rc=$?"
timeout 900 ncu --set full --clock-control none -k regex:"k_materialise|k_refine_finalize|k_finalize|k_member_gemm|k_es_select|k_evolve_batch" -s 168 -c 14 -o gpurun_out/${tag}_full_aux -f $CMD > gpurun_out/${tag}_ncu_full_aux.log 2>&1; echo "ncu full aux rc=$?"
# K7 is launched band by band (3 launches per call at D = 13.4 M): skip 3 calls, capture the 3 bands of the fourth
timeout 600 ncu --set full --clock-control none -k regex:"k_gradient" -s 9 -c 3 -o gpurun_out/${tag}_full_master -f python bench.py --steps 1 --warmup 1 --population 148 --cpu-seconds 1 --nes-pairs 0 > gpurun_out/${tag}_ncu_full_master.log 2>&1; echo "ncu full master rc=$?"
for k in dense aux master; do
  python tools/ncu_full.py gpurun_out/${tag}_full_$k.ncu-rep > gpurun_out/${tag}_full_${k}_summary.txt 2>&1
  [ "$k" != dense ] && rm -f gpurun_out/${tag}_full_$k.ncu-rep
done
grep -E "Kernel Name|gpu__time_duration.sum|dram__bytes_read.sum |dram__bytes_write.sum |pipe_tensor_cycles|lts__throughput|xbar2l1tex_read_bytes.sum " gpurun_out/${tag}_full_dense_summary.txt | head -40
du -sh gpurun_out
