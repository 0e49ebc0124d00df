#!/bin/bash
# call 27 (1 GPU): default bench line of the round's last source (per-level roofline keys from the r2h capture of the same source)
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/c27_bench_full.json 2> gpurun_out/c27_bench_full.err
echo "bench rc=$?" >> gpurun_out/c27_bench_full.err
