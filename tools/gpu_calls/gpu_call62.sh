#!/bin/bash
# two-pipe pass loop A/B (RT_PIPES=1 vs 2) without per-kernel events; new pipes test; bench with the split timed / attribution steps
mkdir -p gpurun_out
L=gpurun_out/pipes_r1j.log; rm -f $L
for p in 1 2; do
  RT_PIPES=$p timeout 600 python tools/pipes_probe.py lamp 3840 2160 640 10 0,2,4,8 >> $L 2>&1
  RT_PIPES=$p timeout 300 python tools/pipes_probe.py lamp 1000 1000 1000 10 0,5 >> $L 2>&1
  RT_PIPES=$p timeout 300 python tools/pipes_probe.py cornell_plus 1000 1000 1000 10 0 >> $L 2>&1
  RT_PIPES=$p timeout 300 python tools/pipes_probe.py teapot 1000 1000 1000 10 0 >> $L 2>&1
done
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "pipes or grouping or split" > gpurun_out/pytest_pipes.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_pipes.log
timeout 900 python bench.py > gpurun_out/bench_r1j.json 2> gpurun_out/bench_r1j.err; echo "bench rc=$?" >> gpurun_out/bench_r1j.err
