#!/bin/bash
mkdir -p gpurun_out
P=gpurun_out/probe16.log
for rep in 1 2; do
echo "== default" >> $P
timeout 600 python tools/perf_probe.py lamp 3840 2160 80 10 group4 noref >> $P 2>&1
echo "== unified pop" >> $P
RT_LIBRTCORE=$PWD/cuda_raytracer_b200/librtcore_upop.so timeout 600 python tools/perf_probe.py lamp 3840 2160 80 10 group4 noref >> $P 2>&1
done
echo "== default 1k" >> $P
timeout 600 python tools/perf_probe.py lamp 1000 1000 320 10 group16 noref >> $P 2>&1
echo "== unified pop 1k" >> $P
RT_LIBRTCORE=$PWD/cuda_raytracer_b200/librtcore_upop.so timeout 600 python tools/perf_probe.py lamp 1000 1000 320 10 group16 noref >> $P 2>&1
RT_LIBRTCORE=$PWD/cuda_raytracer_b200/librtcore_upop.so timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_upop.log 2>&1
