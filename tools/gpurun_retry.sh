#!/bin/bash
# usage: tools/gpurun_retry.sh LOG TIMEOUT [--gpus N] -- 'command'   (retries while the pod answers busy / transient)
LOG=$1; shift; TO=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $TO "$@" > $LOG 2>&1; rc=$?
  if grep -q "status=transient\|nothing was charged" $LOG || [ $rc = 3 ]; then sleep 90; continue; fi
  break
done
echo "rc=$rc try=$i" >> $LOG
