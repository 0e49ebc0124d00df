#!/bin/bash
# call 4 (2 GPUs): rt_multi over NCCL, gpus=2 front end, bench at N=2 (lamp 4K with the image check, teapot)
mkdir -p gpurun_out
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
