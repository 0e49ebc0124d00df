#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_parity.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_parity.log
for lib in ab/librtcore_nospec.so cuda_raytracer_b200/librtcore.so ab/librtcore_nospec.so cuda_raytracer_b200/librtcore.so; do
  echo "== $lib" >> gpurun_out/ab_spec.log
  RT_LIBRTCORE=$PWD/$lib timeout 600 python tools/perf_probe.py lamp 3840 2160 100 10 group5 noref 2>&1 | grep -v "shade ms/bounce\|live per" >> gpurun_out/ab_spec.log
done
for lib in ab/librtcore_nospec.so cuda_raytracer_b200/librtcore.so; do
  echo "== $lib" >> gpurun_out/ab_spec.log
  RT_LIBRTCORE=$PWD/$lib timeout 600 python tools/perf_probe.py teapot 1920 1080 100 10 group5 noref 2>&1 | grep -v "ms/bounce\|live per" >> gpurun_out/ab_spec.log
  RT_LIBRTCORE=$PWD/$lib timeout 600 python tools/perf_probe.py cornell_plus 1000 1000 100 10 group5 noref 2>&1 | grep -v "ms/bounce\|live per" >> gpurun_out/ab_spec.log
done
