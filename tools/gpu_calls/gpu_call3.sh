#!/bin/bash
mkdir -p gpurun_out
export RT_STACK=local
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for knobs in "RT_BATCH=96 RT_REFILL=22" "RT_BATCH=32 RT_REFILL=22" "RT_BATCH=192 RT_REFILL=22" "RT_BATCH=96 RT_REFILL=12" "RT_BATCH=96 RT_REFILL=28" "RT_BATCH=96 RT_REFILL=32"; do
  echo "== $knobs" >> gpurun_out/probe3.log
  env $knobs timeout 600 python tools/perf_probe.py lamp 1000 1000 100 10 nosort noref >> gpurun_out/probe3.log 2>&1
  env $knobs timeout 600 python tools/perf_probe.py lamp 3840 2160 40 10 nosort noref >> gpurun_out/probe3.log 2>&1
done
TCMD="python tools/ncu_target.py lamp 1000 1000"
$TCMD > gpurun_out/plain_target.log 2>&1 && timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'trace_rays_kernel|shade_kernel' -c 20 -o gpurun_out/prof_split $TCMD > gpurun_out/ncu_full.log 2>&1
