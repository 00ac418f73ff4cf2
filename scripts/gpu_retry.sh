#!/bin/bash
# gpu_retry.sh LOGFILE TIMEOUT -- 'command': run gpurun, retrying while the pod answers busy (exit 3, nothing charged).
log=$1; shift; to=$1; shift
for k in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$to" "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
