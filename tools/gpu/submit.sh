#!/bin/bash
# usage: tools/gpu/submit.sh <script-under-tools/gpu> <timeout-seconds> : gpurun with retries while the pod answers busy
S=$1; T=${2:-900}
for i in $(seq 1 12); do
  /usr/local/graft/bin/gpurun --timeout $T -- "bash tools/gpu/$S" > /tmp/${S%.sh}.out 2>&1
  if grep -q "status=transient\|rc=3\|exit code 3" /tmp/${S%.sh}.out; then sleep 45; continue; fi
  break
done
tail -5 /tmp/${S%.sh}.out
