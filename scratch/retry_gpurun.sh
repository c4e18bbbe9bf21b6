#!/bin/bash
# usage: scratch/retry_gpurun.sh GPUS TIMEOUT SCRIPT [ATTEMPTS]   retries while the pod answers busy (exit 3)
GPUS=$1; TMO=$2; SCRIPT=$3; N=${4:-8}
for i in $(seq 1 $N); do
  /usr/local/graft/bin/gpurun --gpus $GPUS --timeout $TMO -- "bash $SCRIPT" > /tmp/retry_gpurun_last.log 2>&1
  rc=$?
  echo "attempt $i rc=$rc" >> /tmp/retry_gpurun.log
  if [ $rc -ne 3 ]; then cat /tmp/retry_gpurun_last.log | tail -40; exit $rc; fi
  sleep 150
done
tail -5 /tmp/retry_gpurun_last.log; exit 3
