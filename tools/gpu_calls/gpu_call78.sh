#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_r1l.json 2> gpurun_out/bench_r1l.err; echo "bench rc=$?" >> gpurun_out/bench_r1l.err
timeout 900 python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_r1l_ref.json 2> gpurun_out/bench_r1l_ref.err; echo "ref rc=$?" >> gpurun_out/bench_r1l_ref.err
for w in lamp_1k teapot_1k spheres_1k; do
  timeout 600 python bench.py --workload $w --no-cpu-baseline --no-ref-gpu > gpurun_out/r1l_bench_$w.json 2> gpurun_out/r1l_bench_$w.err
done
