#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/ab_tile.log
for lib in cuda_raytracer_b200/librtcore.so ab/librtcore_tile2.so ab/librtcore_tile4.so ab/librtcore_tile16.so; do
  echo "== $lib" >> gpurun_out/ab_tile.log
  RT_LIBRTCORE=$PWD/$lib timeout 600 python tools/perf_probe.py lamp 3840 2160 160 10 group8 noref 2>&1 | grep "group\|trace ms" >> gpurun_out/ab_tile.log
  RT_LIBRTCORE=$PWD/$lib timeout 600 python tools/perf_probe.py lamp 1000 1000 320 10 group16 noref 2>&1 | grep "group\|trace ms" >> gpurun_out/ab_tile.log
  RT_LIBRTCORE=$PWD/$lib timeout 600 python tools/perf_probe.py teapot 1920 1080 100 10 group5 noref 2>&1 | grep "group\|trace ms" >> gpurun_out/ab_tile.log
done
