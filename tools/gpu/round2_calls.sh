#!/bin/bash
# Every GPU call of round 2 as it was run (provenance of the numbers under profiles/r2*): `tools/gpu/round2_calls.sh <nn>` runs call nn
# on the box; from here: `bash tools/gpu/run_gpu.sh <log> <timeout s> <gpus> "bash tools/gpu/round2_calls.sh <nn>"` (run_gpu.sh retries
# while the pod is busy). Each call writes into gpurun_out/, which gpurun merges back; the files worth keeping were copied to profiles/.
mkdir -p gpurun_out
case "$1" in
01)  # was tools/gpu/r2_call01_nodes.sh
# call 1: correctness of node modes + ray ranges, then node-format A/B
nvidia-smi -L > gpurun_out/smi.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "compressed or ray_ranges or closest_hit or pass_split or image_vs_oracle or primary_entry" > gpurun_out/c1_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c1_tests.log
timeout 600 python tools/nodes_probe.py lamp 3840 2160 160 10 f32,q16,w4 3 > gpurun_out/c1_nodes_lamp4k.jsonl 2> gpurun_out/c1_nodes_lamp4k.err
timeout 300 python tools/nodes_probe.py teapot 1000 1000 400 10 f32,q16,w4 3 > gpurun_out/c1_nodes_teapot.jsonl 2> gpurun_out/c1_nodes_teapot.err
timeout 300 python tools/nodes_probe.py cornell_plus 1000 1000 400 10 f32,q16,w4 3 > gpurun_out/c1_nodes_cplus.jsonl 2> gpurun_out/c1_nodes_cplus.err
PROBE_BVH=sbvh timeout 600 python tools/nodes_probe.py lamp 3840 2160 160 10 f32,q16 3 > gpurun_out/c1_nodes_lamp4k_sbvh.jsonl 2> gpurun_out/c1_nodes_lamp4k_sbvh.err
echo done
;;
02)  # was tools/gpu/r2_call02.sh
# call 2: whole GPU suite on the new build (ray ranges, node formats, multi handle, N2 upload path), bench lines (N=1) for
# lamp 4K + reference arm, upload timing
nvidia-smi -L > gpurun_out/smi.txt
timeout 1500 python -m pytest tests -x -q -m gpu -s > gpurun_out/c2_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c2_tests.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/c2_bench_lamp4k.json 2> gpurun_out/c2_bench_lamp4k.err
echo "bench rc=$?" >> gpurun_out/c2_bench_lamp4k.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/c2_bench_reference.json 2> gpurun_out/c2_bench_reference.err
echo "ref rc=$?" >> gpurun_out/c2_bench_reference.err
timeout 300 python bench.py --workload teapot_1k --steps 5 --warmup 3 --no-cpu-baseline --no-ref-gpu --no-extras > gpurun_out/c2_bench_teapot_n1.json 2> gpurun_out/c2_bench_teapot_n1.err
python __graft_entry__.py smoke > gpurun_out/c2_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/c2_smoke.log
echo done
;;
03)  # was tools/gpu/r2_call03.sh
# call 3: GPU suite again (sphere loop now contraction-free), radix-sort variants vs CUB, ncu full capture f32 + q16
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
;;
03s)  # was tools/gpu/r2_call03_sort.sh
# call 3: radix-sort digit-pass variants against CUB (tools/sort_probe.py), 16 M and 64 M pairs
for rank in 0 1 2; do for items in 12 16 18; do
  echo "== RT_SORT_RANK=$rank RT_SORT_ITEMS=$items" >> gpurun_out/c3_sort.log
  RT_SORT_RANK=$rank RT_SORT_ITEMS=$items PROBE_NO_CUB=$([ "$rank$items" = "016" ] && echo 0 || echo 1) timeout 300 python tools/sort_probe.py 16e6,64e6 >> gpurun_out/c3_sort.log 2>&1
done; done
timeout 300 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "radix" >> gpurun_out/c3_sort.log 2>&1
echo done
;;
04)  # was tools/gpu/r2_call04_2gpu.sh
# call 4 (2 GPUs): rt_multi over NCCL, gpus=2 front end, bench at N=2 (lamp 4K with the image check, teapot)
nvidia-smi -L > gpurun_out/c4_smi.txt
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_frontend.py -q -m gpu -k "multi_device or gpus_word or deep_leaf or env_map or spheres" > gpurun_out/c4_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c4_tests.log
run() {  # run <n> <workload> <out>
  if [ $1 -eq 1 ]; then timeout 900 python bench.py --gpus 1 --workload $2 --no-cpu-baseline --no-ref-gpu --no-extras > gpurun_out/$3.json 2> gpurun_out/$3.err
  else timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --workload $2 --no-cpu-baseline --no-ref-gpu > gpurun_out/$3.json 2> gpurun_out/$3.err; fi
  echo "n=$1 $2 rc=$?" >> gpurun_out/c4_rc.txt
}
run 2 lamp_4k c4_bench_lamp4k_n2
run 1 teapot_1k c4_bench_teapot_n1
run 2 teapot_1k c4_bench_teapot_n2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > gpurun_out/c4_bench_reference_n2.json 2> gpurun_out/c4_bench_reference_n2.err
echo "ref n=2 rc=$?" >> gpurun_out/c4_rc.txt
echo done
;;
05)  # was tools/gpu/r2_call05.sh
# call 5: parity of the collective leaf phase, then A/B: base (round-1 tracer) / ftz / bf / nols (ftz+bf) / main (ftz+bf+leaf share)
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_bvh_build.py -q -m gpu -x > gpurun_out/c5_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c5_tests.log
bash tools/gpu/ab_probe.sh c5_ab nols base ftz bf
;;
06)  # was tools/gpu/r2_call06.sh
# call 6: A/B of the tracer variants (absolute library paths this time), threshold sweep, sort tile sizes
bash tools/gpu/ab_probe.sh c6_ab base ftz bf ls
bash tools/gpu/tune_sweep.sh c6_tune
for items in 18 22 26; do
  echo "== RT_SORT_RANK=2 RT_SORT_ITEMS=$items" >> gpurun_out/c6_sort.log
  RT_SORT_RANK=2 RT_SORT_ITEMS=$items PROBE_NO_CUB=$([ "$items" = "18" ] && echo 0 || echo 1) timeout 300 python tools/sort_probe.py 16e6,64e6 >> gpurun_out/c6_sort.log 2>&1
done
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -q -m gpu -k "radix" >> gpurun_out/c6_sort.log 2>&1
;;
07)  # was tools/gpu/r2_call07_8gpu.sh
# call 7 (8 GPUs): BASELINE configs[3] = teapot.scene as shipped at 1/2/4/8 GPUs, lamp 4K at 8 (image check), rt_multi over 8
# devices + front-end gpus=2/4/8
nvidia-smi -L > gpurun_out/c7_smi.txt
run() {  # run <n> <workload> <out>
  if [ $1 -eq 1 ]; then timeout 600 python bench.py --gpus 1 --workload $2 --steps 5 --warmup 3 --no-cpu-baseline --no-ref-gpu --no-extras > gpurun_out/$3.json 2> gpurun_out/$3.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --workload $2 --steps 5 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/$3.json 2> gpurun_out/$3.err; fi
  echo "n=$1 $2 rc=$?" >> gpurun_out/c7_rc.txt
}
for n in 1 2 4 8; do run $n teapot_1k c7_bench_teapot_n$n; done
run 8 lamp_4k c7_bench_lamp4k_n8
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_frontend.py -q -m gpu -k "multi_device or gpus_word" > gpurun_out/c7_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c7_tests.log
# the front end over 8 GPUs on the headline frame: `GPU Took` of gpus=8 against gpus=1
cd scenes && for g in 1 8; do
  sed 's/^image .*/image 3840 2160 640 10 0.1/' lamp.scene > /tmp/lamp4k.scene
  timeout 300 ../cuda_raytracer_b200/bin/raytracing /tmp/lamp4k.scene no_sort stats gpus=$g out=/tmp/lamp_g$g.png > ../gpurun_out/c7_frontend_g$g.log 2>&1
  echo "frontend gpus=$g rc=$?" >> ../gpurun_out/c7_rc.txt
done
cd ..
python - <<'PY' > gpurun_out/c7_frontend_cmp.log 2>&1
import sys, numpy as np
sys.path[:0] = ["tests", "oracle", "."]
from util import read_png_rgb8
a, b = read_png_rgb8("/tmp/lamp_g1.png"), read_png_rgb8("/tmp/lamp_g8.png")
print("4K png gpus=1 vs gpus=8: 8-bit RMSE", float(np.sqrt(np.mean((a.astype(float) - b.astype(float)) ** 2))), "max abs", int(np.abs(a.astype(int) - b.astype(int)).max()))
PY
echo done
;;
08)  # was tools/gpu/r2_call08.sh
# call 8 (1 GPU): whole GPU suite on the build with the sample split, smoke
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c8_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c8_tests.log
python __graft_entry__.py smoke > gpurun_out/c8_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/c8_smoke.log
;;
09)  # was tools/gpu/r2_call09.sh
# call 9 (1 GPU): the no-alloc test alone in a fresh process, then the whole suite (it is collected last), then bench N=1
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/c9_noalloc_alone.log 2>&1
echo "alone rc=$?" >> gpurun_out/c9_noalloc_alone.log
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c9_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c9_tests.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/c9_bench_lamp4k.json 2> gpurun_out/c9_bench_lamp4k.err
echo "bench rc=$?" >> gpurun_out/c9_bench_lamp4k.err
;;
10)  # was tools/gpu/r2_call10.sh
# call 10 (1 GPU): tracer occupancy A/B, reorder pass on/off, bench lines of the six 1000x1000 workloads with the reference GPU
# build timed beside them (least-squares rate +- spread)
bash tools/gpu/ab_probe.sh c10_ab ctas8 ctas10
for cfg in "lamp 3840 2160 80 10" "cornell_plus 1000 1000 200 10" "teapot 1000 1000 200 10"; do
  timeout 600 python tools/reorder_probe.py $cfg >> gpurun_out/c10_reorder.jsonl 2>> gpurun_out/c10_reorder.err
done
for w in lamp_1k teapot_1k cornell_1k cornell_plus_1k spheres_1k glass_teapot_1k; do
  timeout 900 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/c10_bench_$w.json 2> gpurun_out/c10_bench_$w.err
  echo "$w rc=$?" >> gpurun_out/c10_rc.txt
done
;;
11)  # was tools/gpu/r2_call11.sh
# call 11 (1 GPU): occupancy A/B of the tracer (10 / 12 CTAs per SM) and of shade_kernel (5 / 6 CTAs per SM, 128-thread CTAs)
bash tools/gpu/ab_probe.sh c11_ab ctas10 ctas12 shade5 shade6 shadeb128
;;
12)  # was tools/gpu/r2_call12.sh
# call 12 (1 GPU): final build: whole GPU suite, ncu --set full of one pass (f32 nodes) for the roofline keys, launch list of the
# bench command, bench line
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c12_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c12_tests.log
timeout 300 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c12_ncu_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'trace_rays_kernel|shade_kernel' -c 8 -f -o gpurun_out/prof_r2e_f32 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c12_ncu_full.log 2>&1
BCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ref-gpu --no-extras"
$BCMD > gpurun_out/c12_bench_plain.json 2> gpurun_out/c12_bench_plain.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2e_launches_bench.csv $BCMD > gpurun_out/c12_ncu_launch.log 2>&1
echo "ncu launch list rc=$?" >> gpurun_out/c12_ncu_launch.log
;;
13)  # was tools/gpu/r2_call13_8gpu.sh
# call 13 (8 GPUs): BASELINE configs[3] (teapot.scene as shipped) and configs[4] (lamp 4K) at 1/2/4/8 GPUs with the per-pixel
# sample split, rt_multi_* tests over NCCL, the front end over 8 GPUs with NCCL and with the peer-access reduce
nvidia-smi -L > gpurun_out/c13_smi.txt
run() {  # run <n> <workload> <out>
  if [ $1 -eq 1 ]; then timeout 600 python bench.py --gpus 1 --workload $2 --steps 3 --warmup 3 --no-cpu-baseline --no-ref-gpu --no-extras > gpurun_out/$3.json 2> gpurun_out/$3.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --workload $2 --steps 3 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/$3.json 2> gpurun_out/$3.err; fi
  echo "n=$1 $2 rc=$?" >> gpurun_out/c13_rc.txt
}
for n in 1 2 4 8; do run $n teapot_1k c13_bench_teapot_n$n; done
for n in 1 2 4 8; do run $n lamp_4k c13_bench_lamp4k_n$n; done
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_frontend.py -q -m gpu -k "multi_device or gpus_word" > gpurun_out/c13_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c13_tests.log
cd scenes
sed 's/^image .*/image 3840 2160 640 10 0.1/' lamp.scene > /tmp/lamp4k.scene
for mode in "gpus=1" "gpus=8" "gpus=8 reduce=p2p"; do
  tag=$(echo $mode | tr ' =' '__')
  timeout 300 ../cuda_raytracer_b200/bin/raytracing /tmp/lamp4k.scene no_sort stats $mode out=/tmp/lamp_$tag.png > ../gpurun_out/c13_frontend_$tag.log 2>&1
  echo "frontend $mode rc=$?" >> ../gpurun_out/c13_rc.txt
done
cd ..
python - <<'PY' > gpurun_out/c13_frontend_cmp.log 2>&1
import sys, numpy as np
sys.path[:0] = ["tests", "oracle", "."]
from util import read_png_rgb8
a = read_png_rgb8("/tmp/lamp_gpus_1.png")
for tag in ("gpus_8", "gpus_8_reduce_p2p"):
    b = read_png_rgb8(f"/tmp/lamp_{tag}.png")
    print("4K png gpus=1 vs", tag, ": 8-bit RMSE", float(np.sqrt(np.mean((a.astype(float) - b.astype(float)) ** 2))), "max abs", int(np.abs(a.astype(int) - b.astype(int)).max()))
PY
echo done
;;
14)  # was tools/gpu/r2_call14.sh
timeout 300 python tools/share_probe.py teapot 1000 1000 1000 10 8 > gpurun_out/c14_share_teapot.jsonl 2>&1
timeout 300 python tools/share_probe.py teapot 1000 1000 1000 10 4 >> gpurun_out/c14_share_teapot.jsonl 2>&1
;;
15)  # was tools/gpu/r2_call15.sh
# call 15 (1 GPU): tail_kernel: parity suite with it forced on (small images: it takes over at bounce 2), then sweeps of the
# threshold on the teapot share (strong scaling at N = 8), teapot, lamp 4K and cornell_plus whole frames
RT_TAIL=2000000 timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_frontend.py -q -m gpu -x > gpurun_out/c15_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c15_tests.log
for t in 0 50000 200000 800000 3000000; do
  echo "== RT_TAIL=$t" >> gpurun_out/c15_share.log
  RT_TAIL=$t timeout 300 python tools/share_probe.py teapot 1000 1000 1000 10 8 2>&1 | grep -E "whole|auto" >> gpurun_out/c15_share.log
  for cfg in "lamp 3840 2160 160 10" "teapot 1000 1000 400 10" "cornell_plus 1000 1000 400 10"; do
    echo "== RT_TAIL=$t $cfg" >> gpurun_out/c15_frames.log
    RT_TAIL=$t timeout 600 python tools/nodes_probe.py $cfg f32 3 >> gpurun_out/c15_frames.log 2>&1
  done
done
;;
16)  # was tools/gpu/r2_call16.sh
RT_TAIL=0 timeout 300 python tools/partial_probe.py > gpurun_out/c16_partial.log 2>&1
;;
17)  # was tools/gpu/r2_call17.sh
# call 17 (1 GPU): partial passes folded into waves + wave sizing rounded up: parity suite, then the per-share cost again
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c17_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c17_tests.log
timeout 300 python tools/partial_probe.py > gpurun_out/c17_partial.log 2>&1
timeout 300 python tools/share_probe.py teapot 1000 1000 1000 10 8 2>&1 | grep -E "whole|auto" > gpurun_out/c17_share.log
timeout 300 python tools/share_probe.py teapot 1000 1000 1000 10 4 2>&1 | grep -E "whole|auto" >> gpurun_out/c17_share.log
for cfg in "lamp 3840 2160 160 10" "teapot 1000 1000 400 10" "cornell_plus 1000 1000 400 10"; do
  timeout 600 python tools/nodes_probe.py $cfg f32 3 >> gpurun_out/c17_frames.log 2>&1
done
;;
18)  # was tools/gpu/r2_call18.sh
# call 18 (1 GPU): where a wave's fixed cost goes
timeout 600 python tools/wave_probe.py > gpurun_out/c18_wave.log 2>&1
echo "rc=$?" >> gpurun_out/c18_wave.log
;;
19)  # was tools/gpu/r2_call19.sh
# call 19 (1 GPU): final build (automatic pipe count): whole GPU suite, shares of ranks 0 / 1 / 7 of 8, ncu --set full of one pass
# (f32 nodes) for the roofline keys, launch list of the bench command, full bench line, whole frames
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c19_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c19_tests.log
timeout 300 python tools/share_probe.py teapot 1000 1000 1000 10 8 0,1,4,7 2>&1 | grep -E "whole|auto" > gpurun_out/c19_share.log
timeout 300 python tools/share_probe.py teapot 1000 1000 1000 10 4 0,1 2>&1 | grep -E "whole|auto" >> gpurun_out/c19_share.log
timeout 300 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c19_ncu_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'trace_rays_kernel|shade_kernel' -c 8 -f -o gpurun_out/prof_r2f_f32 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c19_ncu_full.log 2>&1
BCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ref-gpu --no-extras"
$BCMD > gpurun_out/c19_bench_plain.json 2> gpurun_out/c19_bench_plain.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2f_launches_bench.csv $BCMD > gpurun_out/c19_ncu_launch.log 2>&1
echo "ncu launch list rc=$?" >> gpurun_out/c19_ncu_launch.log
for cfg in "lamp 3840 2160 160 10" "teapot 1000 1000 400 10" "cornell_plus 1000 1000 400 10"; do
  timeout 600 python tools/nodes_probe.py $cfg f32 3 >> gpurun_out/c19_frames.log 2>&1
done
timeout 900 python bench.py > gpurun_out/c19_bench_full.json 2> gpurun_out/c19_bench_full.err
echo "bench rc=$?" >> gpurun_out/c19_bench_full.err
;;
20)  # was tools/gpu/r2_call20_8gpu.sh
# call 20 (8 GPUs): teapot.scene as shipped at 1/2/4/8 GPUs and lamp 4K at 1/8 on the build with partial passes folded into waves and
# the automatic pipe count
run() {  # run <n> <workload> <out>
  if [ $1 -eq 1 ]; then timeout 600 python bench.py --gpus 1 --workload $2 --steps 5 --warmup 3 --no-cpu-baseline --no-ref-gpu --no-extras > gpurun_out/$3.json 2> gpurun_out/$3.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --workload $2 --steps 5 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/$3.json 2> gpurun_out/$3.err; fi
  echo "n=$1 $2 rc=$?" >> gpurun_out/c20_rc.txt
}
for n in 1 8 4 2; do run $n teapot_1k c20_bench_teapot_n$n; done
for n in 1 8; do run $n lamp_4k c20_bench_lamp4k_n$n; done
echo done
;;
21)  # was tools/gpu/r2_call21.sh
# call 21 (1 GPU): A/B of RT_SHADE_SPLIT (hit / miss rays of a CTA's window dealt to separate warps in shade_kernel, bounces >= 1):
# parity tests on the variant, whole frames + per-bounce kernel times of variant and base from the same source
RT_LIBRTCORE=$PWD/ab/split/librtcore.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -q -m gpu -x > gpurun_out/c21_tests_split.log 2>&1
echo "tests rc=$?" >> gpurun_out/c21_tests_split.log
for v in base0 split base0 split; do
  for cfg in "lamp 3840 2160 160 10" "teapot 1000 1000 400 10" "cornell_plus 1000 1000 400 10"; do
    echo "== $v $cfg" >> gpurun_out/c21_frames.log
    RT_LIBRTCORE=$PWD/ab/$v/librtcore.so timeout 600 python tools/nodes_probe.py $cfg f32 3 >> gpurun_out/c21_frames.log 2>&1
  done
done
;;
22)  # was tools/gpu/r2_call22.sh
# call 22 (1 GPU): committed tree: whole GPU suite + smoke, one GPU's share of 8 on every bundled scene as shipped (1000x1000, 1000 spp),
# default bench line
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c22_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c22_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/c22_smoke.log 2>&1
for s in cornell spheres cornell_plus glass_teapot lamp; do
  echo "== $s" >> gpurun_out/c22_share.log
  timeout 300 python tools/share_probe.py $s 1000 1000 1000 10 8 0,1 2>&1 | grep -E "whole|auto" >> gpurun_out/c22_share.log
done
timeout 900 python bench.py > gpurun_out/c22_bench_full.json 2> gpurun_out/c22_bench_full.err
echo "bench rc=$?" >> gpurun_out/c22_bench_full.err
;;
23)  # was tools/gpu/r2_call23.sh
# call 23 (1 GPU): committed tree with source-hash build ids: GPU suite, default bench line (per-level roofline keys from the
# r2f capture of the same sources), reference arm
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c23_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c23_tests.log
timeout 900 python bench.py > gpurun_out/c23_bench_full.json 2> gpurun_out/c23_bench_full.err
echo "bench rc=$?" >> gpurun_out/c23_bench_full.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/c23_bench_ref.json 2> gpurun_out/c23_bench_ref.err
echo "ref rc=$?" >> gpurun_out/c23_bench_ref.err
;;
24)  # was tools/gpu/r2_call24_8gpu.sh
# call 24 (8 GPUs): the other bundled scenes as shipped (1000x1000, 1000 spp) on 8 GPUs; their one-GPU frames on this build are in
# profiles/r2f_share_all_scenes.log ("whole frame") and lamp_1k is run here at N=1 too
run() {  # run <n> <workload> <out>
  if [ $1 -eq 1 ]; then timeout 600 python bench.py --gpus 1 --workload $2 --steps 5 --warmup 3 --no-cpu-baseline --no-ref-gpu --no-extras > gpurun_out/$3.json 2> gpurun_out/$3.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --workload $2 --steps 5 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/$3.json 2> gpurun_out/$3.err; fi
  echo "n=$1 $2 rc=$?" >> gpurun_out/c24_rc.txt
}
run 8 lamp_1k c24_bench_lamp_1k_n8
run 1 lamp_1k c24_bench_lamp_1k_n1
run 8 spheres_1k c24_bench_spheres_1k_n8
run 8 glass_teapot_1k c24_bench_glass_teapot_1k_n8
run 8 cornell_plus_1k c24_bench_cornell_plus_1k_n8
echo done
;;
25)  # was tools/gpu/r2_call25.sh
# call 25 (1 GPU): balanced wave plan: whole GPU suite, shares of 8 (lamp 1000x1000 ranks 0 / 1 / 3, teapot ranks 0 / 1), ncu --set full
# of one pass for the roofline keys of this build, whole frames
timeout 900 python -m pytest tests -q -m gpu -x > gpurun_out/c25_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c25_tests.log
timeout 200 python tools/share_probe.py lamp 1000 1000 1000 10 8 0,1,3 2>&1 | grep -E "whole|auto" > gpurun_out/c25_share.log
timeout 200 python tools/share_probe.py teapot 1000 1000 1000 10 8 0,1 2>&1 | grep -E "whole|auto" >> gpurun_out/c25_share.log
timeout 200 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c25_ncu_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'trace_rays_kernel|shade_kernel' -c 8 -f -o gpurun_out/prof_r2g_f32 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c25_ncu_full.log 2>&1
for cfg in "lamp 3840 2160 160 10" "teapot 1000 1000 1000 10" "cornell_plus 1000 1000 400 10"; do
  timeout 300 python tools/nodes_probe.py $cfg f32 3 >> gpurun_out/c25_frames.log 2>&1
done
;;
26)  # was tools/gpu/r2_call26.sh
# call 26 (1 GPU): final source (first-fit waves for one-round calls, equal waves otherwise): whole GPU suite, ncu --set full of one
# pass for the roofline keys of this build, teapot shares of 8 and whole frames
timeout 900 python -m pytest tests -q -m gpu -x > gpurun_out/c26_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c26_tests.log
timeout 200 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c26_ncu_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'trace_rays_kernel|shade_kernel' -c 8 -f -o gpurun_out/prof_r2h_f32 python tools/ncu_target.py lamp 3840 2160 nosort f32 > gpurun_out/c26_ncu_full.log 2>&1
timeout 200 python tools/share_probe.py teapot 1000 1000 1000 10 8 0,1 2>&1 | grep -E "whole|auto" > gpurun_out/c26_share.log
timeout 200 python tools/nodes_probe.py lamp 3840 2160 160 10 f32 3 >> gpurun_out/c26_frames.log 2>&1
;;
27)  # was tools/gpu/r2_call27.sh
# call 27 (1 GPU): default bench line of the round's last source (per-level roofline keys from the r2h capture of the same source)
timeout 600 python bench.py > gpurun_out/c27_bench_full.json 2> gpurun_out/c27_bench_full.err
echo "bench rc=$?" >> gpurun_out/c27_bench_full.err
;;
28)  # was tools/gpu/r2_call28.sh
# call 28 (1 GPU, last of the round): the pipe-count guard (host side): wave / pipe / range tests and teapot's share of 8 again
timeout 150 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "pipes or sample_ranges or ray_ranges or grouping or pass_split" > gpurun_out/c28_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c28_tests.log
timeout 60 python tools/share_probe.py teapot 1000 1000 1000 10 8 0 2>&1 | grep -E "whole|auto" > gpurun_out/c28_share.log
;;
29)  # whole GPU suite + smoke on the round's last source
timeout 400 python -m pytest tests -q -m gpu -x > gpurun_out/c29_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c29_tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/c29_smoke.log 2>&1
;;
30)  # lamp 1000x1000: per-bounce and per-pass kernel times of rank 0's and rank 3's share of 8 (why rank 3 is 1.7 ms slower)
timeout 150 python tools/share_bounces.py lamp 1000 1000 1000 10 8 0 3 > gpurun_out/c30_share_bounces.log 2>&1
;;
31)  # last source: kernel id on the box (cuobjdump there) and a short bench line whose roofline keys come from the r2h capture
python -c "from cuda_raytracer_b200.core import kernel_id, build_id; print(kernel_id(), build_id())" > gpurun_out/c31_ids.log 2>&1
timeout 200 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ref-gpu --no-extras > gpurun_out/c31_bench.json 2> gpurun_out/c31_bench.err
echo "bench rc=$?" >> gpurun_out/c31_bench.err
;;
32)  # planner moved into wave_plan.cuh behind rt_plan_waves (host-only refactor): whole GPU suite + teapot's share of 8
timeout 300 python -m pytest tests -q -m gpu -x > gpurun_out/c32_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c32_tests.log
timeout 60 python tools/share_probe.py teapot 1000 1000 1000 10 8 0 2>&1 | grep -E "whole|auto" > gpurun_out/c32_share.log
;;
*) echo "usage: $0 <call>: 01 02 03 03s 04 05 06 07 08 09 10 11 12 13 14 15 16 17 18 19 20 21 22 23 24 25 26 27 28 29 30 31 32"; exit 2 ;;
esac
