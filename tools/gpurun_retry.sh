#!/bin/bash
# usage: tools/gpurun_retry.sh <log> <timeout_s> [--gpus N] -- '<command>'   (retries while the pod answers "busy")
LOG=$1; shift; TMO=$1; shift
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun --timeout $TMO "$@" > $LOG 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc (attempt $i)" >> $LOG; exit $rc; fi
  echo "attempt $i: busy" >> ${LOG}.retries
  sleep 90
done
