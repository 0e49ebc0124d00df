#!/bin/bash
mkdir -p gpurun_out
P="python tools/perf_probe.py lamp 3840 2160 100 10 group5 noref"
run() { echo "== $1" >> gpurun_out/ab_tune2.log; env $2 timeout 600 $P 2>&1 | grep "group5" >> gpurun_out/ab_tune2.log; }
run "default" "RT_X=0"
for rf in 8 10 12 14; do run "refill $rf" "RT_REFILL=$rf"; done
for rf in 12 14 16; do for nm in 12 16 20; do run "refill $rf nodemin $nm batch 192" "RT_REFILL=$rf RT_NODEMIN=$nm RT_BATCH=192"; done; done
run "refill 16 nodemin 16 batch 384" "RT_REFILL=16 RT_NODEMIN=16 RT_BATCH=384"
P="python tools/perf_probe.py lamp 1000 1000 320 10 group16 noref"
run "1k default" "RT_X=0"
run "1k refill 16 nodemin 16 batch 192" "RT_REFILL=16 RT_NODEMIN=16 RT_BATCH=192"
run "1k refill 12 nodemin 16 batch 192" "RT_REFILL=12 RT_NODEMIN=16 RT_BATCH=192"
P="python tools/perf_probe.py teapot 1920 1080 100 10 group5 noref"
run "teapot default" "RT_X=0"
run "teapot refill 16 nodemin 16 batch 192" "RT_REFILL=16 RT_NODEMIN=16 RT_BATCH=192"
run "teapot refill 12 nodemin 16 batch 192" "RT_REFILL=12 RT_NODEMIN=16 RT_BATCH=192"
P="python tools/perf_probe.py cornell_plus 1000 1000 100 10 group5 noref"
run "cornell_plus default" "RT_X=0"
run "cornell_plus refill 16 nodemin 16 batch 192" "RT_REFILL=16 RT_NODEMIN=16 RT_BATCH=192"
run "cornell_plus refill 12 nodemin 16 batch 192" "RT_REFILL=12 RT_NODEMIN=16 RT_BATCH=192"
