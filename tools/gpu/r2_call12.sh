#!/bin/bash
# call 12 (1 GPU): final build: whole GPU suite, ncu --set full of one pass (f32 nodes) for the roofline keys, launch list of the
# bench command, bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c12_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c12_tests.log
timeout 300 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c12_ncu_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'trace_rays_kernel|shade_kernel' -c 8 -f -o gpurun_out/prof_r2e_f32 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c12_ncu_full.log 2>&1
BCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ref-gpu --no-extras"
$BCMD > gpurun_out/c12_bench_plain.json 2> gpurun_out/c12_bench_plain.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2e_launches_bench.csv $BCMD > gpurun_out/c12_ncu_launch.log 2>&1
echo "ncu launch list rc=$?" >> gpurun_out/c12_ncu_launch.log
