#!/bin/bash
# usage: scripts/gpuretry.sh <log> <gpus> <timeout> <cmd>   (retries while the pool answers busy; nothing is charged for those)
log=$1; g=$2; to=$3; shift 3
for i in $(seq 1 60); do
  /usr/local/graft/bin/gpurun --gpus $g --timeout $to -- "$@" > $log 2>&1
  if grep -q "status=transient\|rc=None\|busy" $log && ! grep -q "charged=[1-9]" $log; then sleep 150; continue; fi
  break
done
echo DONE >> $log
