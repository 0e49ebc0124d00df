#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/diag_trees.log; rm -f $L
echo "== default (CH)" >> $L
timeout 600 python tools/diag_trees.py cornell cornell_plus teapot >> $L 2>&1
echo "== ab/top6 (lo/hi boxes)" >> $L
RT_LIBRTCORE=$PWD/ab/top6/librtcore.so timeout 600 python tools/diag_trees.py cornell cornell_plus teapot >> $L 2>&1
timeout 900 python -m pytest tests/test_gpu_bvh_build.py -m gpu -q > gpurun_out/pytest_bvh.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_bvh.log
