#!/bin/bash
# call 3: radix-sort digit-pass variants against CUB (tools/sort_probe.py), 16 M and 64 M pairs
mkdir -p gpurun_out
for rank in 0 1 2; do for items in 12 16 18; do
  echo "== RT_SORT_RANK=$rank RT_SORT_ITEMS=$items" >> gpurun_out/c3_sort.log
  RT_SORT_RANK=$rank RT_SORT_ITEMS=$items PROBE_NO_CUB=$([ "$rank$items" = "016" ] && echo 0 || echo 1) timeout 300 python tools/sort_probe.py 16e6,64e6 >> gpurun_out/c3_sort.log 2>&1
done; done
timeout 300 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "radix" >> gpurun_out/c3_sort.log 2>&1
echo done
