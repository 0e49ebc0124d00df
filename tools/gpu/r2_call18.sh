#!/bin/bash
# call 18 (1 GPU): where a wave's fixed cost goes
mkdir -p gpurun_out
timeout 600 python tools/wave_probe.py > gpurun_out/c18_wave.log 2>&1
echo "rc=$?" >> gpurun_out/c18_wave.log
