#!/bin/bash
mkdir -p gpurun_out
P="python tools/perf_probe.py lamp 3840 2160 100 10 group5 noref"
run() { echo "== $1" >> gpurun_out/ab_tune.log; env $2 timeout 600 $P 2>&1 | grep "group5" >> gpurun_out/ab_tune.log; }
run "default" "RT_X=0"
for top in 4 6 12; do run "stack top $top" "RT_LIBRTCORE=$PWD/ab/librtcore_top$top.so"; done
run "default again" "RT_X=0"
for rf in 16 19 25 28; do run "refill $rf" "RT_REFILL=$rf"; done
for nm in 4 6 10 12 16; do run "nodemin $nm" "RT_NODEMIN=$nm"; done
run "refill 25 nodemin 10" "RT_REFILL=25 RT_NODEMIN=10"
run "refill 19 nodemin 6" "RT_REFILL=19 RT_NODEMIN=6"
run "batch 64" "RT_BATCH=64"
run "batch 192" "RT_BATCH=192"
