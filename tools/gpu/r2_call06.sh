#!/bin/bash
# call 6: A/B of the tracer variants (absolute library paths this time), threshold sweep, sort tile sizes
mkdir -p gpurun_out
bash tools/gpu/ab_probe.sh c6_ab base ftz bf ls
bash tools/gpu/tune_sweep.sh c6_tune
for items in 18 22 26; do
  echo "== RT_SORT_RANK=2 RT_SORT_ITEMS=$items" >> gpurun_out/c6_sort.log
  RT_SORT_RANK=2 RT_SORT_ITEMS=$items PROBE_NO_CUB=$([ "$items" = "18" ] && echo 0 || echo 1) timeout 300 python tools/sort_probe.py 16e6,64e6 >> gpurun_out/c6_sort.log 2>&1
done
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -q -m gpu -k "radix" >> gpurun_out/c6_sort.log 2>&1
