#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for sc in "lamp 1000 1000 100 10" "lamp 3840 2160 40 10" "teapot 1000 1000 100 10" "cornell 1000 1000 100 10"; do
  timeout 600 python tools/perf_probe.py $sc nosort noref >> gpurun_out/probe6.log 2>&1
done
