#!/bin/bash
# call 22 (1 GPU): committed tree: whole GPU suite + smoke, one GPU's share of 8 on every bundled scene as shipped (1000x1000, 1000 spp),
# default bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c22_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c22_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/c22_smoke.log 2>&1
for s in cornell spheres cornell_plus glass_teapot lamp; do
  echo "== $s" >> gpurun_out/c22_share.log
  timeout 300 python tools/share_probe.py $s 1000 1000 1000 10 8 0,1 2>&1 | grep -E "whole|auto" >> gpurun_out/c22_share.log
done
timeout 900 python bench.py > gpurun_out/c22_bench_full.json 2> gpurun_out/c22_bench_full.err
echo "bench rc=$?" >> gpurun_out/c22_bench_full.err
