#!/bin/bash
# usage: tools/gpurun_retry.sh <timeout> <logname> [--gpus N] -- <command>   (retries while no box / slot is free)
T=$1; LOG=$2; shift 2
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout $T "$@" > /tmp/gpurun_$LOG.txt 2>&1
  rc=$?
  if grep -q "status=transient\|nothing was charged" /tmp/gpurun_$LOG.txt && [ $rc -ne 0 ]; then sleep 45; continue; fi
  break
done
tail -60 /tmp/gpurun_$LOG.txt
exit $rc
