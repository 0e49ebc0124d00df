#!/bin/bash
# call 28 (1 GPU, last of the round): the pipe-count guard (host side): wave / pipe / range tests and teapot's share of 8 again
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "pipes or sample_ranges or ray_ranges or grouping or pass_split" > gpurun_out/c28_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c28_tests.log
timeout 60 python tools/share_probe.py teapot 1000 1000 1000 10 8 0 2>&1 | grep -E "whole|auto" > gpurun_out/c28_share.log
