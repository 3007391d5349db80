#!/bin/bash
# experiment: A/B two builds of the library on one box (A = the in-tree build, B = nes-img-captioning_b200/libnes_b200_$1.so), A B A B
d=nes-img-captioning_b200
cp $d/libnes_b200.so /tmp/libA.so
for leg in A B A B; do
  if [ $leg = A ]; then cp /tmp/libA.so $d/libnes_b200.so; else cp $d/libnes_b200_$1.so $d/libnes_b200.so; fi
  python bench.py --steps 10 --warmup 3 --no-extras --nes-pairs 0 2>/dev/null > gpurun_out/ablib_$leg.json; python tools/show_bench.py gpurun_out/ablib_$leg.json $leg
done
cp /tmp/libA.so $d/libnes_b200.so
