#!/bin/bash
# call 25 (1 GPU): balanced wave plan: whole GPU suite, shares of 8 (lamp 1000x1000 ranks 0 / 1 / 3, teapot ranks 0 / 1), ncu --set full
# of one pass for the roofline keys of this build, whole frames
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x > gpurun_out/c25_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c25_tests.log
timeout 200 python tools/share_probe.py lamp 1000 1000 1000 10 8 0,1,3 2>&1 | grep -E "whole|auto" > gpurun_out/c25_share.log
timeout 200 python tools/share_probe.py teapot 1000 1000 1000 10 8 0,1 2>&1 | grep -E "whole|auto" >> gpurun_out/c25_share.log
timeout 200 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c25_ncu_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'trace_rays_kernel|shade_kernel' -c 8 -f -o gpurun_out/prof_r2g_f32 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c25_ncu_full.log 2>&1
for cfg in "lamp 3840 2160 160 10" "teapot 1000 1000 1000 10" "cornell_plus 1000 1000 400 10"; do
  timeout 300 python tools/nodes_probe.py $cfg f32 3 >> gpurun_out/c25_frames.log 2>&1
done
