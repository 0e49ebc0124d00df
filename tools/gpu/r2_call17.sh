#!/bin/bash
# call 17 (1 GPU): partial passes folded into waves + wave sizing rounded up: parity suite, then the per-share cost again
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c17_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c17_tests.log
timeout 300 python tools/partial_probe.py > gpurun_out/c17_partial.log 2>&1
timeout 300 python tools/share_probe.py teapot 1000 1000 1000 10 8 2>&1 | grep -E "whole|auto" > gpurun_out/c17_share.log
timeout 300 python tools/share_probe.py teapot 1000 1000 1000 10 4 2>&1 | grep -E "whole|auto" >> gpurun_out/c17_share.log
for cfg in "lamp 3840 2160 160 10" "teapot 1000 1000 400 10" "cornell_plus 1000 1000 400 10"; do
  timeout 600 python tools/nodes_probe.py $cfg f32 3 >> gpurun_out/c17_frames.log 2>&1
done
