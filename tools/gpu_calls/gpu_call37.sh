#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -x -q -k "sort" > gpurun_out/pytest_sort.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_sort.log
timeout 600 python tools/sort_probe.py 16e6,64e6 > gpurun_out/sort_probe2.log 2>&1; echo "rc=$?" >> gpurun_out/sort_probe2.log
