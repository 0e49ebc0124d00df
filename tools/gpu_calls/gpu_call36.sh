#!/bin/bash
mkdir -p gpurun_out
TCMD="python tools/sort_probe.py 64e6"
$TCMD > gpurun_out/plain_sort.log 2>&1 && timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'sort_scatter_kernel|sort_histogram_kernel|Onesweep|Histogram' -c 14 -f -o gpurun_out/prof_sort $TCMD > gpurun_out/ncu_sort.log 2>&1
echo "ncu rc=$?" >> gpurun_out/ncu_sort.log
