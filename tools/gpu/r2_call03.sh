#!/bin/bash
# call 3: GPU suite again (sphere loop now contraction-free), radix-sort variants vs CUB, ncu full capture f32 + q16
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -s > gpurun_out/c3_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c3_tests.log
bash tools/gpu/r2_call03_sort.sh
python - <<'PY' > gpurun_out/c3_l2probe.log 2>&1
import cuda_raytracer_b200 as rt
r = rt.Renderer(0)
for mb in (4, 8, 16, 32, 48, 256, 2048):
    print(mb, "MB x2:", round(r.probe_copy_gbs(mb << 20, 400 if mb <= 48 else 20), 1), "GB/s")
PY
# ncu: the plain runs first (each must exit 0), then ONE ncu session per variant in the same call
timeout 300 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c3_ncu_plain_f32.log 2>&1 && \
timeout 300 python tools/ncu_target.py lamp 3840 2160 nosort q16 > gpurun_out/c3_ncu_plain_q16.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'trace_rays_kernel|shade_kernel' -c 8 -f -o gpurun_out/prof_r2a_f32 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c3_ncu_f32.log 2>&1
timeout 900 ncu --set full --clock-control none -k regex:'trace_rays_kernel' -c 4 -f -o gpurun_out/prof_r2a_q16 python tools/ncu_target.py lamp 3840 2160 nosort q16 > gpurun_out/c3_ncu_q16.log 2>&1
echo done
