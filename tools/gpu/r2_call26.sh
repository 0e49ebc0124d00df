#!/bin/bash
# call 26 (1 GPU): final source (first-fit waves for one-round calls, equal waves otherwise): whole GPU suite, ncu --set full of one
# pass for the roofline keys of this build, teapot shares of 8 and whole frames
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x > gpurun_out/c26_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c26_tests.log
timeout 200 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c26_ncu_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'trace_rays_kernel|shade_kernel' -c 8 -f -o gpurun_out/prof_r2h_f32 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c26_ncu_full.log 2>&1
timeout 200 python tools/share_probe.py teapot 1000 1000 1000 10 8 0,1 2>&1 | grep -E "whole|auto" > gpurun_out/c26_share.log
timeout 200 python tools/nodes_probe.py lamp 3840 2160 160 10 f32 3 >> gpurun_out/c26_frames.log 2>&1
