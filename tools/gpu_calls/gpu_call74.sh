#!/bin/bash
# A/B: shade_kernel with more registers / fewer CTAs per SM (launch-bounds minimum 1, 2, 3) against the default
mkdir -p gpurun_out
L=gpurun_out/shade_occ2_r1k.log; rm -f $L
for v in base sh1 sh2 sh3; do
  lib=$PWD/ab/$v/librtcore.so; [ $v = base ] && lib=$PWD/cuda_raytracer_b200/librtcore.so
  echo "== variant $v" >> $L
  RT_LIBRTCORE=$lib timeout 600 python tools/pipes_probe.py lamp 3840 2160 640 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 300 python tools/pipes_probe.py teapot 1000 1000 1000 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 300 python tools/pipes_probe.py cornell_plus 1000 1000 1000 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 600 python tools/perf_probe.py lamp 3840 2160 160 10 group8 noref 2>&1 | grep "group\|shade ms/bounce" >> $L
done
