#!/bin/bash
# call 19 (1 GPU): final build (automatic pipe count): whole GPU suite, shares of ranks 0 / 1 / 7 of 8, ncu --set full of one pass
# (f32 nodes) for the roofline keys, launch list of the bench command, full bench line, whole frames
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c19_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c19_tests.log
timeout 300 python tools/share_probe.py teapot 1000 1000 1000 10 8 0,1,4,7 2>&1 | grep -E "whole|auto" > gpurun_out/c19_share.log
timeout 300 python tools/share_probe.py teapot 1000 1000 1000 10 4 0,1 2>&1 | grep -E "whole|auto" >> gpurun_out/c19_share.log
timeout 300 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c19_ncu_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'trace_rays_kernel|shade_kernel' -c 8 -f -o gpurun_out/prof_r2f_f32 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c19_ncu_full.log 2>&1
BCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ref-gpu --no-extras"
$BCMD > gpurun_out/c19_bench_plain.json 2> gpurun_out/c19_bench_plain.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2f_launches_bench.csv $BCMD > gpurun_out/c19_ncu_launch.log 2>&1
echo "ncu launch list rc=$?" >> gpurun_out/c19_ncu_launch.log
for cfg in "lamp 3840 2160 160 10" "teapot 1000 1000 400 10" "cornell_plus 1000 1000 400 10"; do
  timeout 600 python tools/nodes_probe.py $cfg f32 3 >> gpurun_out/c19_frames.log 2>&1
done
timeout 900 python bench.py > gpurun_out/c19_bench_full.json 2> gpurun_out/c19_bench_full.err
echo "bench rc=$?" >> gpurun_out/c19_bench_full.err
