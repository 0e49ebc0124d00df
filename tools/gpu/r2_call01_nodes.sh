#!/bin/bash
# call 1: correctness of node modes + ray ranges, then node-format A/B
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "compressed or ray_ranges or closest_hit or pass_split or image_vs_oracle or primary_entry" > gpurun_out/c1_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c1_tests.log
timeout 600 python tools/nodes_probe.py lamp 3840 2160 160 10 f32,q16,w4 3 > gpurun_out/c1_nodes_lamp4k.jsonl 2> gpurun_out/c1_nodes_lamp4k.err
timeout 300 python tools/nodes_probe.py teapot 1000 1000 400 10 f32,q16,w4 3 > gpurun_out/c1_nodes_teapot.jsonl 2> gpurun_out/c1_nodes_teapot.err
timeout 300 python tools/nodes_probe.py cornell_plus 1000 1000 400 10 f32,q16,w4 3 > gpurun_out/c1_nodes_cplus.jsonl 2> gpurun_out/c1_nodes_cplus.err
PROBE_BVH=sbvh timeout 600 python tools/nodes_probe.py lamp 3840 2160 160 10 f32,q16 3 > gpurun_out/c1_nodes_lamp4k_sbvh.jsonl 2> gpurun_out/c1_nodes_lamp4k_sbvh.err
echo done
