#!/bin/bash
# rays claimed per atomic by a tracer warp (RT_BATCH), default build, two pipes
mkdir -p gpurun_out
L=gpurun_out/batch_r1m.log; rm -f $L
run() { echo "== $1" >> $L; env $1 timeout 300 python tools/pipes_probe.py lamp 3840 2160 640 10 0 2>&1 | grep group >> $L; env $1 timeout 300 python tools/pipes_probe.py teapot 1000 1000 1000 10 0 2>&1 | grep group >> $L; env $1 timeout 300 python tools/pipes_probe.py cornell_plus 1000 1000 1000 10 0 2>&1 | grep group >> $L; }
for b in 192 256 384 512 768 192; do run "RT_BATCH=$b"; done
