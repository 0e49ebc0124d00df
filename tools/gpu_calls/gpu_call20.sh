#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_bvh_build.py -m gpu -x -q -s > gpurun_out/pytest_bvh.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_bvh.log
RT_BUILD_TRACE=1 timeout 300 python tools/build_probe.py lamp 2>&1 | tail -40 > gpurun_out/build_probe.log
for big in 512 1024 4096 8192; do echo "== RT_BUILD_BIG=$big" >> gpurun_out/build_probe2.log; RT_BUILD_BIG=$big timeout 300 python tools/build_probe.py teapot lamp >> gpurun_out/build_probe2.log 2>&1; done
