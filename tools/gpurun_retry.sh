#!/bin/bash
# Retry a gpurun call while the pod answers "busy" (exit 3) or drains; everything else is returned as is.
# usage: tools/gpurun_retry.sh <log> <timeout-seconds> <command> [gpus]
log=$1; to=$2; cmd=$3; gpus=${4:-1}
for attempt in 1 2 3 4 5 6 7 8; do
  if [ "$gpus" = "1" ]; then
    /usr/local/graft/bin/gpurun --timeout "$to" -- "$cmd" > "$log" 2>&1
  else
    /usr/local/graft/bin/gpurun --gpus "$gpus" --timeout "$to" -- "$cmd" > "$log" 2>&1
  fi
  rc=$?
  if [ $rc -ne 3 ] && ! grep -qi "draining" "$log"; then
    echo "gpurun rc=$rc (attempt $attempt)"; exit $rc
  fi
  sleep 90
done
echo "gpurun gave up"; exit 3
