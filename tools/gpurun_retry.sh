#!/bin/bash
# usage: tools/gpurun_retry.sh <log> <timeout-seconds> <command> [gpus]
# retries while the pool answers "busy / draining" (exit 3); stops on anything else
log=$1; to=$2; cmd=$3; gpus=${4:-1}
extra=""; [ "$gpus" != "1" ] && extra="--gpus $gpus"
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun $extra --timeout "$to" -- "$cmd" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc (attempt $i)" >> "$log"; exit $rc; fi
  sleep 45
done
echo "gpurun: gave up after 40 busy answers" >> "$log"; exit 3
