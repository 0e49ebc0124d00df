#!/bin/bash
# tracer thresholds re-swept on the current build (runtime-tunable variant ab/tune), two pipes, whole-loop time
mkdir -p gpurun_out
L=gpurun_out/tune_r1m.log; rm -f $L
run() { echo "== $1" >> $L; env $1 RT_LIBRTCORE=$PWD/ab/tune/librtcore.so timeout 300 python tools/pipes_probe.py lamp 3840 2160 320 10 0 2>&1 | grep group >> $L; }
run "RT_X=0"
for rf in 12 16 20 24; do for nm in 8 12 16; do run "RT_REFILL=$rf RT_NODEMIN=$nm"; done; done
run "RT_BATCH=96"
run "RT_BATCH=384"
