#!/bin/bash
# A/B builds of the render core with extra -D flags: tools/build_variant.sh <name> [-DRT_X=1 ...] -> ab/<name>/librtcore.so
# (same ABI; select with RT_LIBRTCORE=ab/<name>/librtcore.so). ab/ is git-ignored and travels with gpurun.
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
name=$1; shift
mkdir -p "$ROOT/ab/$name"
CS=$ROOT/cuda_raytracer_b200/csrc
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -ccbin /usr/bin/g++ -Xcompiler -fPIC "$@" \
    -c "$CS/device/rt_core.cu" -o "$ROOT/ab/$name/rt_core.o"
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared --cudart static -ccbin /usr/bin/g++ \
    -o "$ROOT/ab/$name/librtcore.so" "$CS"/_obj/scene_loader.o "$CS"/_obj/bvh_builder.o "$CS"/_obj/png_writer.o "$CS"/_obj/scene_api.o "$ROOT/ab/$name/rt_core.o" -lz
echo "built ab/$name/librtcore.so"
