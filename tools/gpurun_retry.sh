#!/bin/bash
# usage: tools/gpurun_retry.sh <timeout-seconds> '<command>'  — retries while the pod answers "transient" (nothing is charged for those)
t=$1; shift
for i in $(seq 1 12); do
    /usr/local/graft/bin/gpurun --timeout "$t" -- "$@" > /tmp/gpurun_retry.$$ 2>&1
    if ! grep -q "status=transient" /tmp/gpurun_retry.$$; then break; fi
    sleep 150
done
cat /tmp/gpurun_retry.$$; rm -f /tmp/gpurun_retry.$$
