#!/bin/bash
# gpu_retry_n.sh LOGFILE TIMEOUT GPUS -- 'command': like gpu_retry.sh with --gpus N
log=$1; shift; to=$1; shift; n=$1; shift
for k in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --gpus "$n" --timeout "$to" "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
