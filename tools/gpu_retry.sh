#!/bin/bash
# tools/gpu_retry.sh <log> <timeout-seconds> '<command>' [gpus]: runs one gpurun call, retrying while the pod is busy (rc 3).
LOG=$1; TO=$2; CMD=$3; G=${4:-1}
for i in $(seq 1 40); do
  if [ "$G" = "1" ]; then /usr/local/graft/bin/gpurun --timeout $TO -- "$CMD" > $LOG 2>&1; else /usr/local/graft/bin/gpurun --gpus $G --timeout $TO -- "$CMD" > $LOG 2>&1; fi
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc (attempt $i)" >> $LOG; exit $rc; fi
  sleep 45
done
echo "gave up" >> $LOG; exit 3
