#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/ab_sb.log
for lib in cuda_raytracer_b200/librtcore.so ab/librtcore_sb128.so ab/librtcore_sb512.so ab/librtcore_sb1024.so; do
  for sc in "lamp 3840 2160 160 10 group8" "cornell_plus 1000 1000 100 10 group5" "teapot 1920 1080 100 10 group5"; do
  echo "== $sc $lib" >> gpurun_out/ab_sb.log
  RT_LIBRTCORE=$PWD/$lib timeout 600 python tools/perf_probe.py $sc noref 2>&1 | grep "group\|ms/bounce" >> gpurun_out/ab_sb.log
  done
done
RT_LIBRTCORE=$PWD/ab/librtcore_sb1024.so timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "image or tail or entry or pass" > gpurun_out/pytest_sb.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_sb.log
