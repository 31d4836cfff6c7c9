#!/bin/bash
# profiles/gpuretry.sh <logfile> <timeout> <command...> : retry a gpurun call while the pod answers "transient"/busy
log=$1; shift; to=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $to -- "$@" > $log 2>&1
  if grep -q "status=transient\|rc=3\|answers busy\|no box" $log; then sleep 90; else break; fi
done
