#!/bin/bash
# A/B: FMA slab / fast rcp / sign masks / constant thresholds / 9 CTAs per SM, + parity tests of the new default build
mkdir -p gpurun_out
L=gpurun_out/ab_r1j.log; rm -f $L
for v in base fma rcp mask opt opt9; do
  for sc in "lamp 3840 2160 160 10 group8"; do
    echo "== $sc $v" >> $L
    RT_LIBRTCORE=$PWD/ab/$v/librtcore.so timeout 600 python tools/perf_probe.py $sc noref 2>&1 | grep "group\|ms/bounce" >> $L
  done
done
for v in base opt; do
  for sc in "cornell_plus 1000 1000 100 10 group5" "teapot 1920 1080 100 10 group5"; do
    echo "== $sc $v" >> $L
    RT_LIBRTCORE=$PWD/ab/$v/librtcore.so timeout 600 python tools/perf_probe.py $sc noref 2>&1 | grep "group\|ms/bounce" >> $L
  done
done
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_bvh_build.py -m gpu -q > gpurun_out/pytest_r1j.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_r1j.log
