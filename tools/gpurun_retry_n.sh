#!/bin/bash
# usage: tools/gpurun_retry_n.sh <gpus> <timeout-seconds> <log> <command...>: gpurun --gpus N, retried while the pod answers busy
N=$1; T=$2; LOG=$3; shift 3
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --gpus $N --timeout $T -- "$@" > $LOG 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
