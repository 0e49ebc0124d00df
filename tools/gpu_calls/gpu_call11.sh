#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 900 python bench.py --no-ref-gpu > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
BCMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-ref-gpu"
$BCMD > gpurun_out/plain_bench.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_bench.csv $BCMD > gpurun_out/ncu_bench.log 2>&1
TCMD="python tools/ncu_target.py lamp 3840 2160"
$TCMD > gpurun_out/plain_target.log 2>&1 && timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'trace_rays_kernel|shade_kernel' -c 8 -o gpurun_out/prof_r1c $TCMD > gpurun_out/ncu_full.log 2>&1
