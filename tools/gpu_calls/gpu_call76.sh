#!/bin/bash
# A/B: multiply-and-shift index division (default build) against the previous build (ab/top6 is NOT comparable: use perf probe of both)
mkdir -p gpurun_out
L=gpurun_out/fastdiv_r1l.log; rm -f $L
timeout 600 python tools/pipes_probe.py lamp 3840 2160 640 10 0 >> $L 2>&1
timeout 300 python tools/pipes_probe.py teapot 1000 1000 1000 10 0 >> $L 2>&1
timeout 300 python tools/pipes_probe.py cornell_plus 1000 1000 1000 10 0 >> $L 2>&1
timeout 300 python tools/pipes_probe.py spheres 1000 1000 1000 10 0 >> $L 2>&1
timeout 600 python tools/perf_probe.py lamp 3840 2160 160 10 group8 noref 2>&1 | grep "group\|ms/bounce" >> $L
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_r1l_full.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_r1l_full.log
