#!/bin/bash
# A/B: shade_kernel with survivors staged in shared memory and the reservation's round trip hidden (ab/stage) against the default
mkdir -p gpurun_out
L=gpurun_out/stage_r1l.log; rm -f $L
for v in base stage; do
  lib=$PWD/ab/$v/librtcore.so; [ $v = base ] && lib=$PWD/cuda_raytracer_b200/librtcore.so
  echo "== variant $v" >> $L
  RT_LIBRTCORE=$lib timeout 600 python tools/pipes_probe.py lamp 3840 2160 640 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 300 python tools/pipes_probe.py teapot 1000 1000 1000 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 300 python tools/pipes_probe.py cornell_plus 1000 1000 1000 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 600 python tools/perf_probe.py lamp 3840 2160 160 10 group8 noref 2>&1 | grep "group\|ms/bounce" >> $L
done
RT_LIBRTCORE=$PWD/ab/stage/librtcore.so timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -q -x > gpurun_out/pytest_stage.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_stage.log
