#!/bin/bash
# usage: tools/gpurun_retry.sh <logfile> <timeout> <command...>  -- retries while the pod answers "transient"/busy
log=$1; shift; to=$1; shift
for attempt in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout $to -- "$@" > $log 2>&1
  if grep -q "status=transient\|rc=3\b\|busy" $log && ! grep -q "status=ok" $log; then sleep 90; continue; fi
  break
done
