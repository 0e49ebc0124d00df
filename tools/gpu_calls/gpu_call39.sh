#!/bin/bash
mkdir -p gpurun_out
for wl in lamp_1k teapot_1k cornell_1k cornell_plus_1k spheres_1k glass_teapot_1k; do
  timeout 600 python bench.py --workload $wl --no-cpu-baseline > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo "$wl rc=$?" >> gpurun_out/bench_all_rc.txt
done
