#!/bin/bash
# call 23 (1 GPU): committed tree with source-hash build ids: GPU suite, default bench line (per-level roofline keys from the
# r2f capture of the same sources), reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c23_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c23_tests.log
timeout 900 python bench.py > gpurun_out/c23_bench_full.json 2> gpurun_out/c23_bench_full.err
echo "bench rc=$?" >> gpurun_out/c23_bench_full.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/c23_bench_ref.json 2> gpurun_out/c23_bench_ref.err
echo "ref rc=$?" >> gpurun_out/c23_bench_ref.err
