#!/bin/bash
# usage: grun.sh <timeout> <gpus> <command...> : gpurun, retried while the pod answers busy
T=$1; G=$2; shift 2
for i in 1 2 3 4 5 6 7 8 9 10 11 12; do
  if [ "$G" = "1" ]; then out=$(/usr/local/graft/bin/gpurun --timeout $T -- "$@" 2>&1); else out=$(/usr/local/graft/bin/gpurun --gpus $G --timeout $T -- "$@" 2>&1); fi
  echo "$out" | tail -40
  if echo "$out" | grep -q "status=transient\|retry in a few minutes\|no box\|busy"; then sleep 120; continue; fi
  break
done
