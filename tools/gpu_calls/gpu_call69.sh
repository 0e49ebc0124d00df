#!/bin/bash
# full ncu capture of the bounce kernels of one 4K pass (bounces 0-3) + launch list of the bench command
mkdir -p gpurun_out
timeout 300 python tools/ncu_target.py lamp 3840 2160 > gpurun_out/ncu_target_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'trace_rays_kernel|shade_kernel' -c 8 -f -o gpurun_out/prof_r1k python tools/ncu_target.py lamp 3840 2160 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?" >> gpurun_out/ncu_full.log
BCMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-ref-gpu"
$BCMD > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err && timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r1k_launches_bench.csv $BCMD > gpurun_out/ncu_launch.log 2>&1
echo "ncu rc=$?" >> gpurun_out/ncu_launch.log
