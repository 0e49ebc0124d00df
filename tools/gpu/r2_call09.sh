#!/bin/bash
# call 9 (1 GPU): the no-alloc test alone in a fresh process, then the whole suite (it is collected last), then bench N=1
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/c9_noalloc_alone.log 2>&1
echo "alone rc=$?" >> gpurun_out/c9_noalloc_alone.log
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c9_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c9_tests.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/c9_bench_lamp4k.json 2> gpurun_out/c9_bench_lamp4k.err
echo "bench rc=$?" >> gpurun_out/c9_bench_lamp4k.err
