#!/bin/bash
mkdir -p gpurun_out
TCMD="python tools/ncu_target.py lamp 3840 2160"
$TCMD > gpurun_out/plain_target.log 2>&1 && timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'trace_rays_kernel|shade_kernel' -c 8 -f -o gpurun_out/prof_r1h $TCMD > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?" >> gpurun_out/ncu_full.log
