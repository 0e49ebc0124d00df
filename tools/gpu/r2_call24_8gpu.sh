#!/bin/bash
# call 24 (8 GPUs): the other bundled scenes as shipped (1000x1000, 1000 spp) on 8 GPUs; their one-GPU frames on this build are in
# profiles/r2f_share_all_scenes.log ("whole frame") and lamp_1k is run here at N=1 too
mkdir -p gpurun_out
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
