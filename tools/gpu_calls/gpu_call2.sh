#!/bin/bash
# Iteration call: parity tests + perf probe of both stack variants.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for mode in smem local; do
  echo "== RT_STACK=$mode" >> gpurun_out/probe.log
  RT_STACK=$mode timeout 600 python tools/perf_probe.py lamp 1000 1000 100 10 nosort,sort15 noref >> gpurun_out/probe.log 2>&1
  RT_STACK=$mode timeout 600 python tools/perf_probe.py lamp 3840 2160 40 10 nosort noref >> gpurun_out/probe.log 2>&1
  RT_STACK=$mode timeout 600 python tools/perf_probe.py teapot 1000 1000 100 10 nosort noref >> gpurun_out/probe.log 2>&1
  RT_STACK=$mode timeout 600 python tools/perf_probe.py cornell 1000 1000 100 10 nosort noref >> gpurun_out/probe.log 2>&1
done
