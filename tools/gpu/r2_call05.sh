#!/bin/bash
# call 5: parity of the collective leaf phase, then A/B: base (round-1 tracer) / ftz / bf / nols (ftz+bf) / main (ftz+bf+leaf share)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_bvh_build.py -q -m gpu -x > gpurun_out/c5_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c5_tests.log
bash tools/gpu/ab_probe.sh c5_ab nols base ftz bf
