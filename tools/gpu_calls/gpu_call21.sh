#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_bvh_build.py tests/test_gpu_frontend.py -m gpu -x -q -s > gpurun_out/pytest_bvh.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_bvh.log
RT_BUILD_TRACE=1 timeout 300 python tools/build_probe.py lamp 2>&1 | tail -8 > gpurun_out/build_probe.log
timeout 300 python tools/build_probe.py teapot lamp >> gpurun_out/build_probe.log 2>&1
for v in "" devbvh "" devbvh; do
  echo "== $v" >> gpurun_out/ab_order.log
  timeout 600 python tools/perf_probe.py lamp 3840 2160 100 10 group5 noref $v 2>&1 | grep -v "ms/bounce\|live per" >> gpurun_out/ab_order.log
done
for v in "" devbvh; do
  echo "== $v" >> gpurun_out/ab_order.log
  timeout 600 python tools/perf_probe.py teapot 1920 1080 100 10 group5 noref $v 2>&1 | grep -v "ms/bounce\|live per" >> gpurun_out/ab_order.log
done
