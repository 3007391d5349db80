#!/bin/bash
# experiment: A/B/C... several builds of the library on ONE box (box-to-box clocks differ by +-8 %): tools/ab_libs.sh "<bench args>" tag1 tag2 ...
# runs nes-img-captioning_b200/libnes_b200_<tag>.so for every tag, the whole list twice; "cur" = the in-tree build
d=nes-img-captioning_b200
args=$1; shift
cp $d/libnes_b200.so /tmp/libcur.so
mkdir -p gpurun_out
for rep in 1 2; do
  for tag in "$@"; do
    if [ $tag = cur ]; then cp /tmp/libcur.so $d/libnes_b200.so; else cp $d/libnes_b200_$tag.so $d/libnes_b200.so; fi
    timeout 600 python bench.py --steps 10 --warmup 3 --no-extras --nes-pairs 0 $args 2> gpurun_out/abl_${tag}_$rep.err > gpurun_out/abl_${tag}_$rep.json
    python tools/show_bench.py gpurun_out/abl_${tag}_$rep.json $tag
  done
done
cp /tmp/libcur.so $d/libnes_b200.so
