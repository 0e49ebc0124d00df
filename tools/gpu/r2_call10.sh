#!/bin/bash
# call 10 (1 GPU): tracer occupancy A/B, reorder pass on/off, bench lines of the six 1000x1000 workloads with the reference GPU
# build timed beside them (least-squares rate +- spread)
mkdir -p gpurun_out
bash tools/gpu/ab_probe.sh c10_ab ctas8 ctas10
for cfg in "lamp 3840 2160 80 10" "cornell_plus 1000 1000 200 10" "teapot 1000 1000 200 10"; do
  timeout 600 python tools/reorder_probe.py $cfg >> gpurun_out/c10_reorder.jsonl 2>> gpurun_out/c10_reorder.err
done
for w in lamp_1k teapot_1k cornell_1k cornell_plus_1k spheres_1k glass_teapot_1k; do
  timeout 900 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/c10_bench_$w.json 2> gpurun_out/c10_bench_$w.err
  echo "$w rc=$?" >> gpurun_out/c10_rc.txt
done
