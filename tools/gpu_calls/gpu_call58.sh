#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/ab_tune3.log
P="python tools/perf_probe.py lamp 3840 2160 160 10 group8 noref"
run() { echo "== $1" >> gpurun_out/ab_tune3.log; env $2 timeout 600 $P 2>&1 | grep "group8" >> gpurun_out/ab_tune3.log; }
run "default 16/12/192" "RT_X=0"
for rf in 14 18 20; do for nm in 10 12 14; do run "refill $rf nodemin $nm" "RT_REFILL=$rf RT_NODEMIN=$nm"; done; done
run "refill 16 nodemin 10" "RT_REFILL=16 RT_NODEMIN=10"
run "refill 16 nodemin 14" "RT_REFILL=16 RT_NODEMIN=14"
run "batch 128" "RT_BATCH=128"
run "batch 256" "RT_BATCH=256"
