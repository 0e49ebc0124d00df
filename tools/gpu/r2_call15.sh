#!/bin/bash
# call 15 (1 GPU): tail_kernel: parity suite with it forced on (small images: it takes over at bounce 2), then sweeps of the
# threshold on the teapot share (strong scaling at N = 8), teapot, lamp 4K and cornell_plus whole frames
mkdir -p gpurun_out
RT_TAIL=2000000 timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_frontend.py -q -m gpu -x > gpurun_out/c15_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c15_tests.log
for t in 0 50000 200000 800000 3000000; do
  echo "== RT_TAIL=$t" >> gpurun_out/c15_share.log
  RT_TAIL=$t timeout 300 python tools/share_probe.py teapot 1000 1000 1000 10 8 2>&1 | grep -E "whole|auto" >> gpurun_out/c15_share.log
  for cfg in "lamp 3840 2160 160 10" "teapot 1000 1000 400 10" "cornell_plus 1000 1000 400 10"; do
    echo "== RT_TAIL=$t $cfg" >> gpurun_out/c15_frames.log
    RT_TAIL=$t timeout 600 python tools/nodes_probe.py $cfg f32 3 >> gpurun_out/c15_frames.log 2>&1
  done
done
