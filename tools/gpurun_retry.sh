#!/bin/bash
# gpurun with retries while the pod answers "busy" (nothing is charged for those): gpurun_retry.sh <timeout> <gpus> <command...>
t=$1; g=$2; shift 2
for i in $(seq 1 40); do
  out=$(/usr/local/graft/bin/gpurun --timeout $t $( [ "$g" != 1 ] && echo --gpus $g ) -- "$@" 2>&1)
  if echo "$out" | grep -q "status=transient"; then sleep 45; continue; fi
  echo "$out" | tail -60
  exit 0
done
echo "gpurun: still busy after 40 tries"
