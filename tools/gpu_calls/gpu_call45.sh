#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -x -q > gpurun_out/pytest_parity.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_parity.log
rm -f gpurun_out/ab_entry.log
for v in "rootstart" "" ; do
  echo "== $v" >> gpurun_out/ab_entry.log
  timeout 600 python tools/perf_probe.py lamp 3840 2160 160 10 group8 noref $v 2>&1 | grep "group\|trace ms" >> gpurun_out/ab_entry.log
done
for k in 2 4; do
  echo "== K=$k" >> gpurun_out/ab_entry.log
  RT_ENTRY_K=$k timeout 600 python tools/perf_probe.py lamp 3840 2160 160 10 group8 noref 2>&1 | grep "group\|trace ms" >> gpurun_out/ab_entry.log
done
for sc in "teapot 1920 1080 100 10 group5" "cornell_plus 1000 1000 100 10 group5" "lamp 1000 1000 320 10 group16"; do
  for v in "rootstart" ""; do
  echo "== $sc $v" >> gpurun_out/ab_entry.log
  timeout 600 python tools/perf_probe.py $sc noref $v 2>&1 | grep "group\|trace ms" >> gpurun_out/ab_entry.log
  done
done
