#!/bin/bash
mkdir -p gpurun_out
export RT_STACK=local
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for knobs in "RT_NODEMIN=0" "RT_NODEMIN=4" "RT_NODEMIN=8" "RT_NODEMIN=12" "RT_NODEMIN=16" "RT_NODEMIN=8 RT_REFILL=28"; do
  echo "== $knobs" >> gpurun_out/probe4.log
  env $knobs timeout 600 python tools/perf_probe.py lamp 1000 1000 100 10 nosort noref >> gpurun_out/probe4.log 2>&1
  env $knobs timeout 600 python tools/perf_probe.py lamp 3840 2160 40 10 nosort noref >> gpurun_out/probe4.log 2>&1
done
