#!/bin/bash
# call 8 (1 GPU): whole GPU suite on the build with the sample split, smoke
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c8_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c8_tests.log
python __graft_entry__.py smoke > gpurun_out/c8_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/c8_smoke.log
