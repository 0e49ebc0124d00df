#!/bin/bash
# A/B: float4 accumulation buffer with 16-byte reductions (default) against three 4-byte atomics (ab/fb3)
mkdir -p gpurun_out
L=gpurun_out/fbv4_r1k.log; rm -f $L
for v in fb3 v4; do
  lib=$PWD/ab/$v/librtcore.so; [ $v = v4 ] && lib=$PWD/cuda_raytracer_b200/librtcore.so
  echo "== variant $v" >> $L
  RT_LIBRTCORE=$lib timeout 600 python tools/pipes_probe.py lamp 3840 2160 640 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 300 python tools/pipes_probe.py teapot 1000 1000 1000 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 300 python tools/pipes_probe.py cornell_plus 1000 1000 1000 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 300 python tools/pipes_probe.py spheres 1000 1000 1000 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 600 python tools/perf_probe.py lamp 3840 2160 160 10 group8 noref 2>&1 | grep "group\|ms/bounce" >> $L
done
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_r1k_full.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_r1k_full.log
