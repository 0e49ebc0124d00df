#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/ab_oct.log
RT_LIBRTCORE=$PWD/ab/librtcore_oct.so timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_parity_oct.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_parity_oct.log
for sc in "lamp 3840 2160 160 10 group8" "lamp 1000 1000 320 10 group16" "teapot 1920 1080 100 10 group5" "cornell_plus 1000 1000 100 10 group5"; do
  for lib in cuda_raytracer_b200/librtcore.so ab/librtcore_oct.so; do
  echo "== $sc $lib" >> gpurun_out/ab_oct.log
  RT_LIBRTCORE=$PWD/$lib timeout 600 python tools/perf_probe.py $sc noref 2>&1 | grep "group\|ms/bounce" >> gpurun_out/ab_oct.log
  done
done
