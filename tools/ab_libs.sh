#!/bin/bash
# A/B two builds of libllie.so on ONE box: gpurun_ab/libllie_{old,new}.so, alternating, same command.
#   bash tools/ab_libs.sh "python bench.py --workload c3 --steps 10 --no-e2e --no-cpu --no-also"
set -u
L=low_light_image_enhancement_b200/libllie.so
for rep in 1 2; do
  for v in old new; do
    cp gpurun_ab/libllie_$v.so $L
    echo "== $v (rep $rep)"
    eval "$1"
  done
done
cp gpurun_ab/libllie_new.so $L
