#!/bin/bash
# usage: tools/gpurun_retry.sh <logfile> <gpurun args...>   re-submits while the pod answers "busy" (exit 3: nothing charged)
log=$1; shift
for i in $(seq 1 60); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 45
done
exit 3
