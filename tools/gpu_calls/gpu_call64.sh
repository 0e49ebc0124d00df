#!/bin/bash
# A/B: centre / half-extent boxes (RT_BOX_CH=1) against the default build; parity suite on the variant
mkdir -p gpurun_out
L=gpurun_out/ch_r1k.log; rm -f $L
for v in base ch; do
  lib=$PWD/ab/$v/librtcore.so; [ $v = base ] && lib=$PWD/cuda_raytracer_b200/librtcore.so
  echo "== variant $v" >> $L
  RT_LIBRTCORE=$lib timeout 600 python tools/pipes_probe.py lamp 3840 2160 640 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 300 python tools/pipes_probe.py teapot 1000 1000 1000 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 300 python tools/pipes_probe.py cornell_plus 1000 1000 1000 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 300 python tools/pipes_probe.py spheres 1000 1000 1000 10 0 >> $L 2>&1
  RT_LIBRTCORE=$lib timeout 600 python tools/perf_probe.py lamp 3840 2160 160 10 group8 noref 2>&1 | grep "N_int\|group\|ms/bounce" >> $L
done
RT_LIBRTCORE=$PWD/ab/ch/librtcore.so timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_bvh_build.py -m gpu -q > gpurun_out/pytest_ch.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_ch.log
