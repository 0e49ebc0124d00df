#!/bin/bash
# study: the experimental spatial-split tree (RT_SBVH=1, host builder) against the reference's tree, same render core
mkdir -p gpurun_out
L=gpurun_out/sbvh_r1m.log; rm -f $L
for s in 0 1; do
  echo "== RT_SBVH=$s" >> $L
  RT_SBVH=$s timeout 300 python tools/pipes_probe.py lamp 3840 2160 320 10 0 2>&1 | grep group >> $L
  RT_SBVH=$s timeout 300 python tools/pipes_probe.py teapot 1000 1000 1000 10 0 2>&1 | grep group >> $L
  RT_SBVH=$s timeout 300 python tools/pipes_probe.py glass_teapot 1000 1000 1000 10 0 2>&1 | grep group >> $L
  RT_SBVH=$s timeout 300 python tools/perf_probe.py lamp 3840 2160 160 10 group8 noref 2>&1 | grep "N_int\|group8" >> $L
done
