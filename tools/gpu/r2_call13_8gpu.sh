#!/bin/bash
# call 13 (8 GPUs): BASELINE configs[3] (teapot.scene as shipped) and configs[4] (lamp 4K) at 1/2/4/8 GPUs with the per-pixel
# sample split, rt_multi_* tests over NCCL, the front end over 8 GPUs with NCCL and with the peer-access reduce
mkdir -p gpurun_out
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
