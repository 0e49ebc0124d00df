#!/bin/bash
# call 2: whole GPU suite on the new build (ray ranges, node formats, multi handle, N2 upload path), bench lines (N=1) for
# lamp 4K + reference arm, upload timing
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi.txt
timeout 1500 python -m pytest tests -x -q -m gpu -s > gpurun_out/c2_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c2_tests.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/c2_bench_lamp4k.json 2> gpurun_out/c2_bench_lamp4k.err
echo "bench rc=$?" >> gpurun_out/c2_bench_lamp4k.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/c2_bench_reference.json 2> gpurun_out/c2_bench_reference.err
echo "ref rc=$?" >> gpurun_out/c2_bench_reference.err
timeout 300 python bench.py --workload teapot_1k --steps 5 --warmup 3 --no-cpu-baseline --no-ref-gpu --no-extras > gpurun_out/c2_bench_teapot_n1.json 2> gpurun_out/c2_bench_teapot_n1.err
python __graft_entry__.py smoke > gpurun_out/c2_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/c2_smoke.log
echo done
