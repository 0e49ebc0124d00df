#!/bin/bash
# call 21 (1 GPU): A/B of RT_SHADE_SPLIT (hit / miss rays of a CTA's window dealt to separate warps in shade_kernel, bounces >= 1):
# parity tests on the variant, whole frames + per-bounce kernel times of variant and base from the same source
mkdir -p gpurun_out
RT_LIBRTCORE=$PWD/ab/split/librtcore.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -q -m gpu -x > gpurun_out/c21_tests_split.log 2>&1
echo "tests rc=$?" >> gpurun_out/c21_tests_split.log
for v in base0 split base0 split; do
  for cfg in "lamp 3840 2160 160 10" "teapot 1000 1000 400 10" "cornell_plus 1000 1000 400 10"; do
    echo "== $v $cfg" >> gpurun_out/c21_frames.log
    RT_LIBRTCORE=$PWD/ab/$v/librtcore.so timeout 600 python tools/nodes_probe.py $cfg f32 3 >> gpurun_out/c21_frames.log 2>&1
  done
done
