#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for sc in "lamp 1000 1000 100 10" "lamp 3840 2160 40 10" "teapot 1000 1000 100 10" "cornell 1000 1000 100 10" "cornell_plus 1000 1000 100 10" "spheres 1000 1000 100 10" "glass_teapot 1000 1000 100 10"; do
  timeout 600 python tools/perf_probe.py $sc nosort,sort15 noref >> gpurun_out/probe5.log 2>&1
done
TCMD="python tools/ncu_target.py lamp 1000 1000"
$TCMD > gpurun_out/plain_target.log 2>&1 && timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'trace_rays_kernel|shade_kernel' -c 8 -o gpurun_out/prof_split2 $TCMD > gpurun_out/ncu_full.log 2>&1
