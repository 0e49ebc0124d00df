#!/bin/bash
# co-residency of the two pipes: tracer grid below full occupancy so the other wave's shade CTAs find room
mkdir -p gpurun_out
L=gpurun_out/cores_r1k.log; rm -f $L
run() { echo "== $1" >> $L; env $1 timeout 600 python tools/pipes_probe.py lamp 3840 2160 640 10 0 2>&1 | grep group >> $L; env $1 timeout 300 python tools/pipes_probe.py teapot 1000 1000 1000 10 0 2>&1 | grep group >> $L; }
run "RT_X=0"
run "RT_TRACE_PER_SM=7"
run "RT_TRACE_PER_SM=6"
run "RT_TRACE_PER_SM=7 RT_SHADE_PER_SM=4"
run "RT_TRACE_PER_SM=7 RT_SHADE_PER_SM=2"
run "RT_TRACE_PER_SM=6 RT_SHADE_PER_SM=4"
run "RT_SHADE_PER_SM=4"
run "RT_PIPES=1 RT_TRACE_PER_SM=7"
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_r1k_full.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_r1k_full.log
