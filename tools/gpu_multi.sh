#!/bin/bash
# usage: gpu_multi.sh "1 2 4 8" — scaling runs of bench.py on one box (launched exactly as the driver does)
LIST=${1:-"1 2"}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi_multi.txt
for n in $LIST; do
  if [ $n -eq 1 ]; then
    timeout 900 python bench.py --gpus 1 --no-cpu-baseline --no-ref-gpu > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --no-cpu-baseline --no-ref-gpu > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
  fi
  echo "n=$n rc=$?" >> gpurun_out/scale_rc.txt
done
