#!/bin/bash
# call 7 (8 GPUs): BASELINE configs[3] = teapot.scene as shipped at 1/2/4/8 GPUs, lamp 4K at 8 (image check), rt_multi over 8
# devices + front-end gpus=2/4/8
mkdir -p gpurun_out
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
