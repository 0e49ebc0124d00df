#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
P=gpurun_out/probe12.log
echo "== default (8 CTAs/SM)" >> $P
timeout 600 python tools/perf_probe.py lamp 3840 2160 80 10 group4 noref >> $P 2>&1
timeout 600 python tools/perf_probe.py lamp 1000 1000 320 10 group16 noref >> $P 2>&1
for n in 10 12; do
  echo "== RT_TRACE_CTAS=$n" >> $P
  RT_LIBRTCORE=$PWD/cuda_raytracer_b200/librtcore_ctas$n.so timeout 600 python tools/perf_probe.py lamp 3840 2160 80 10 group4 noref >> $P 2>&1
  RT_LIBRTCORE=$PWD/cuda_raytracer_b200/librtcore_ctas$n.so timeout 600 python tools/perf_probe.py lamp 1000 1000 320 10 group16 noref >> $P 2>&1
done
echo "== smem stack" >> $P
RT_STACK=smem timeout 600 python tools/perf_probe.py lamp 3840 2160 80 10 group4 noref >> $P 2>&1
