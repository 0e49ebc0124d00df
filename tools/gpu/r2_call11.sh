#!/bin/bash
# call 11 (1 GPU): occupancy A/B of the tracer (10 / 12 CTAs per SM) and of shade_kernel (5 / 6 CTAs per SM, 128-thread CTAs)
mkdir -p gpurun_out
bash tools/gpu/ab_probe.sh c11_ab ctas10 ctas12 shade5 shade6 shadeb128
