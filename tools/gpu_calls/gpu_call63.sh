#!/bin/bash
# pipes 2/3/4; trace CTAs per SM and shared stack-top size under two pipes
mkdir -p gpurun_out
L=gpurun_out/pipes_r1k.log; rm -f $L
for p in 2 3 4; do
  RT_PIPES=$p timeout 600 python tools/pipes_probe.py lamp 3840 2160 640 10 0,1,2 >> $L 2>&1
  RT_PIPES=$p timeout 300 python tools/pipes_probe.py lamp 1000 1000 1000 10 0 >> $L 2>&1
  RT_PIPES=$p timeout 300 python tools/pipes_probe.py teapot 1000 1000 1000 10 0 >> $L 2>&1
done
for v in ctas9 top6 top12; do
  echo "== variant $v" >> $L
  RT_LIBRTCORE=$PWD/ab/$v/librtcore.so timeout 600 python tools/pipes_probe.py lamp 3840 2160 640 10 0 >> $L 2>&1
  RT_LIBRTCORE=$PWD/ab/$v/librtcore.so timeout 300 python tools/pipes_probe.py teapot 1000 1000 1000 10 0 >> $L 2>&1
done
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "pipes or grouping or split" > gpurun_out/pytest_pipes.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_pipes.log
