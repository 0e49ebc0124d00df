#!/bin/bash
# bench lines of the other bundled scenes as shipped (parity-test cases, not the headline), r1k build
mkdir -p gpurun_out
for w in lamp_1k teapot_1k cornell_1k cornell_plus_1k spheres_1k glass_teapot_1k; do
  timeout 600 python bench.py --workload $w --no-cpu-baseline --no-ref-gpu > gpurun_out/r1k_bench_$w.json 2> gpurun_out/r1k_bench_$w.err; echo "$w rc=$?" >> gpurun_out/r1k_bench_rc.txt
done
