#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python tools/perf_probe.py lamp 1000 1000 320 10 group1,group2,group4,group8,group16 noref >> gpurun_out/probe7.log 2>&1
timeout 600 python tools/perf_probe.py lamp 3840 2160 160 10 group1,group2,group4,group8 noref >> gpurun_out/probe7.log 2>&1
timeout 600 python tools/perf_probe.py cornell 1000 1000 160 10 group1,group4,group8 noref >> gpurun_out/probe7.log 2>&1
