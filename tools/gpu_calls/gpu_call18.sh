#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sort" > gpurun_out/pytest_sort.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_sort.log
timeout 600 python tools/sort_probe.py 1e6,16e6,64e6 > gpurun_out/sort_probe.log 2>&1; echo "rc=$?" >> gpurun_out/sort_probe.log
timeout 600 python tools/perf_probe.py lamp 3840 2160 100 10 nosort,sort15,group5 noref > gpurun_out/probe_lamp4k_sort.log 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_r1d.json 2> gpurun_out/bench_r1d.err; echo "bench rc=$?" >> gpurun_out/bench_r1d.err
