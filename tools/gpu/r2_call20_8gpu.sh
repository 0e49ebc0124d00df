#!/bin/bash
# call 20 (8 GPUs): teapot.scene as shipped at 1/2/4/8 GPUs and lamp 4K at 1/8 on the build with partial passes folded into waves and
# the automatic pipe count
mkdir -p gpurun_out
run() {  # run <n> <workload> <out>
  if [ $1 -eq 1 ]; then timeout 600 python bench.py --gpus 1 --workload $2 --steps 5 --warmup 3 --no-cpu-baseline --no-ref-gpu --no-extras > gpurun_out/$3.json 2> gpurun_out/$3.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --workload $2 --steps 5 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/$3.json 2> gpurun_out/$3.err; fi
  echo "n=$1 $2 rc=$?" >> gpurun_out/c20_rc.txt
}
for n in 1 8 4 2; do run $n teapot_1k c20_bench_teapot_n$n; done
for n in 1 8; do run $n lamp_4k c20_bench_lamp4k_n$n; done
echo done
