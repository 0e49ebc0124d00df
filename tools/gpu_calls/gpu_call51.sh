#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_parity.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_parity.log
rm -f gpurun_out/ab_tail.log
for sc in "lamp 3840 2160 160 10 group8" "lamp 1000 1000 320 10 group16" "teapot 1920 1080 100 10 group5" "cornell_plus 1000 1000 100 10 group5" "glass_teapot 1000 1000 320 10 group16"; do
  for v in "notail" ""; do
  echo "== $sc $v" >> gpurun_out/ab_tail.log
  timeout 600 python tools/perf_probe.py $sc noref $v 2>&1 | grep "group\|trace ms" >> gpurun_out/ab_tail.log
  done
done
