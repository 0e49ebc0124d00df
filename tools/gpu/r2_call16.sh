#!/bin/bash
mkdir -p gpurun_out
RT_TAIL=0 timeout 300 python tools/partial_probe.py > gpurun_out/c16_partial.log 2>&1
