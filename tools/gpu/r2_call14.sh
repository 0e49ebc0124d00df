#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/share_probe.py teapot 1000 1000 1000 10 8 > gpurun_out/c14_share_teapot.jsonl 2>&1
timeout 300 python tools/share_probe.py teapot 1000 1000 1000 10 4 >> gpurun_out/c14_share_teapot.jsonl 2>&1
