#!/bin/bash
# usage: ab_probe.sh <tag> <variant names...>  — nodes_probe (f32 nodes) of lamp 4K 160 spp and teapot for every ab/<name>/librtcore.so + the main build
tag=$1; shift
mkdir -p gpurun_out
for v in main "$@"; do
  lib=""; [ "$v" != "main" ] && lib="$PWD/ab/$v/librtcore.so"
  for cfg in "lamp 3840 2160 160 10" "teapot 1000 1000 400 10" "cornell_plus 1000 1000 400 10"; do
    echo "== $v $cfg" >> gpurun_out/${tag}.log
    RT_LIBRTCORE=$lib timeout 600 python tools/nodes_probe.py $cfg f32 3 >> gpurun_out/${tag}.log 2>&1
  done
done
echo done
