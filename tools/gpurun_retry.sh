#!/bin/bash
# usage: tools/gpurun_retry.sh LOG TIMEOUT [--gpus N] -- 'command'   retries while the pod answers busy (exit 3)
LOG=$1; shift; TO=$1; shift
for attempt in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout $TO "$@" > $LOG 2>&1; rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" $LOG; then echo "done rc=$rc attempt=$attempt" >> $LOG; exit $rc; fi
  sleep 120
done
echo "gave up" >> $LOG; exit 3
