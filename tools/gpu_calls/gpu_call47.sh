#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/ab_greedy.log
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "entry or image_vs_oracle" > gpurun_out/pytest_parity.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_parity.log
for k in 7 5 3; do
  echo "== K=$k" >> gpurun_out/ab_greedy.log
  RT_ENTRY_K=$k timeout 600 python tools/perf_probe.py lamp 3840 2160 160 10 group8 noref 2>&1 | grep "group\|trace ms" >> gpurun_out/ab_greedy.log
done
timeout 600 python tools/perf_probe.py lamp 1000 1000 320 10 group16 noref 2>&1 | grep "group\|trace ms" >> gpurun_out/ab_greedy.log
timeout 600 python tools/perf_probe.py teapot 1920 1080 100 10 group5 noref 2>&1 | grep "group\|trace ms" >> gpurun_out/ab_greedy.log
