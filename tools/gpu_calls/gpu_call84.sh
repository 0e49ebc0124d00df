#!/bin/bash
# last small sweeps on the final build: tracer CTAs per SM in the grid (occupancy gives 9), wave size 3 vs auto
mkdir -p gpurun_out
L=gpurun_out/last_r1m.log; rm -f $L
run() { echo "== $1" >> $L; env $1 timeout 300 python tools/pipes_probe.py lamp 3840 2160 640 10 $2 2>&1 | grep group >> $L; env $1 timeout 300 python tools/pipes_probe.py teapot 1000 1000 1000 10 0 2>&1 | grep group >> $L; }
run "RT_X=0" 0,3,5
run "RT_TRACE_PER_SM=8" 0
run "RT_BATCH=1024" 0
