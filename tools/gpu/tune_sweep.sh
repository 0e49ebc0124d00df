#!/bin/bash
# usage: tune_sweep.sh <tag>  — tracer thresholds (descent exit / refill) with the RT_TUNE_RUNTIME build, lamp 4K 80 spp
tag=$1
mkdir -p gpurun_out
for nm in 6 10 12 16 20; do for rf in 12 16 20 24; do
  echo "== nodemin $nm refill $rf" >> gpurun_out/${tag}.log
  RT_LIBRTCORE=$PWD/ab/tune/librtcore.so RT_NODEMIN=$nm RT_REFILL=$rf timeout 300 python tools/nodes_probe.py lamp 3840 2160 80 10 f32 2 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: print(l.strip()); continue
    print(round(d['mrays_per_s']), d['trace_ms_per_bounce'][:4])" >> gpurun_out/${tag}.log
done; done
