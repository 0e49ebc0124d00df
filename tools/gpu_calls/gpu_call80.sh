#!/bin/bash
# A/B: e2.z of a leaf's triangles shared through the first record (ab/e2z4) against the default; parity on the variant
mkdir -p gpurun_out
L=gpurun_out/e2z4_r1m.log; rm -f $L
for v in base e2z4; do
  lib=$PWD/ab/$v/librtcore.so; [ $v = base ] && lib=$PWD/cuda_raytracer_b200/librtcore.so
  echo "== variant $v" >> $L
  RT_LIBRTCORE=$lib timeout 600 python tools/pipes_probe.py lamp 3840 2160 640 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 300 python tools/pipes_probe.py teapot 1000 1000 1000 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 300 python tools/pipes_probe.py cornell_plus 1000 1000 1000 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 600 python tools/perf_probe.py lamp 3840 2160 160 10 group8 noref 2>&1 | grep "group\|trace ms/bounce" >> $L
done
RT_LIBRTCORE=$PWD/ab/e2z4/librtcore.so timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_bvh_build.py -m gpu -q -x > gpurun_out/pytest_e2z4.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_e2z4.log
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_bvh_build.py -m gpu -q -x > gpurun_out/pytest_base.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_base.log
