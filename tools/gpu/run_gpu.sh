#!/bin/bash
# usage: run_gpu.sh <logfile> <timeout> <gpus> <command...>  — retries while the pod answers busy (exit 3)
log=$1; to=$2; gpus=$3; shift 3
for attempt in 1 2 3 4 5 6 7 8 9 10; do
  if [ "$gpus" = "1" ]; then /usr/local/graft/bin/gpurun --timeout $to -- "$@" > $log 2>&1; rc=$?
  else /usr/local/graft/bin/gpurun --gpus $gpus --timeout $to -- "$@" > $log 2>&1; rc=$?; fi
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc attempt=$attempt" >> $log; exit $rc; fi
  sleep 90
done
echo "gave up" >> $log
