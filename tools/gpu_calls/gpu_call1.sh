#!/bin/bash
# First GPU call of a session: parity tests, bench (both arms), launch list, full capture of the bounce kernel.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 600 python bench.py --workload lamp_1k --sort 1 --no-cpu-baseline --no-ref-gpu > gpurun_out/bench_lamp1k_sort.json 2>> gpurun_out/bench.err
timeout 600 python bench.py --workload lamp_1k --sort 0 --no-cpu-baseline --no-ref-gpu > gpurun_out/bench_lamp1k_nosort.json 2>> gpurun_out/bench.err
BCMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-ref-gpu"
$BCMD > gpurun_out/plain_bench.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches_bench.csv $BCMD > gpurun_out/ncu_bench.log 2>&1
TCMD="python tools/ncu_target.py lamp 1000 1000"
$TCMD > gpurun_out/plain_target.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:bounce_kernel -c 10 -o gpurun_out/prof_bounce $TCMD > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
