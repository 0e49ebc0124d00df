#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 900 python bench.py --no-ref-gpu --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 600 python tools/perf_probe.py lamp 3840 2160 80 10 group4 noref > gpurun_out/probe14.log 2>&1
timeout 600 python tools/perf_probe.py lamp 1000 1000 320 10 group16 noref >> gpurun_out/probe14.log 2>&1
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 7 python tools/ncu_target.py cornell_plus 96 64 sort > gpurun_out/sanitizer_memcheck.log 2>&1; echo "memcheck rc=$?" >> gpurun_out/sanitizer_memcheck.log
