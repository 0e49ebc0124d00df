#!/bin/bash
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_r1i.json 2> gpurun_out/bench_r1i.err; echo "bench rc=$?" >> gpurun_out/bench_r1i.err
BCMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-ref-gpu"
$BCMD > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err && timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r1i_launches_bench.csv $BCMD > gpurun_out/ncu_launch.log 2>&1
echo "ncu rc=$?" >> gpurun_out/ncu_launch.log
