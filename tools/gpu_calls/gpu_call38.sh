#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/ab_sort.log
for lib in cuda_raytracer_b200/librtcore.so ab/librtcore_base3.so ab/librtcore_noatomic.so ab/librtcore_ctas3.so ab/librtcore_nohist4.so; do
  echo "== $lib" >> gpurun_out/ab_sort.log
  RT_LIBRTCORE=$PWD/$lib timeout 600 python tools/sort_probe.py 64e6 2>&1 | grep "core" >> gpurun_out/ab_sort.log
done
