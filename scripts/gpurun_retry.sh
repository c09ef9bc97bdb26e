#!/bin/bash
# Retry a gpurun call until the pod takes it (status "transient" = nothing charged).  usage: gpurun_retry.sh LOG [gpurun args...]
log=$1; shift
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  if ! grep -q "status=transient" "$log"; then exit 0; fi
  sleep 45
done
