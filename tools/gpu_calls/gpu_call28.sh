#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python tools/perf_probe.py lamp 1000 1000 320 10 group16 noref 2>&1 | grep "group16" > gpurun_out/probe_1k.log
RT_REFILL=22 RT_NODEMIN=8 RT_BATCH=96 timeout 600 python tools/perf_probe.py lamp 1000 1000 320 10 group16 noref 2>&1 | grep "group16" >> gpurun_out/probe_1k.log
timeout 900 python bench.py > gpurun_out/bench_r1f.json 2> gpurun_out/bench_r1f.err; echo "bench rc=$?" >> gpurun_out/bench_r1f.err
