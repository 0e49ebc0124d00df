#!/bin/bash
mkdir -p gpurun_out
RT_BUILD_TRACE=1 timeout 300 python tools/build_probe.py lamp > gpurun_out/build_probe.log 2>&1
for lib in ab/librtcore_old.so cuda_raytracer_b200/librtcore.so ab/librtcore_old.so cuda_raytracer_b200/librtcore.so; do
  echo "== $lib" >> gpurun_out/ab_stack.log
  RT_LIBRTCORE=$PWD/$lib timeout 600 python tools/perf_probe.py lamp 3840 2160 100 10 group5 noref 2>&1 | grep -v "ms/bounce\|live per" >> gpurun_out/ab_stack.log
done
