#!/bin/bash
# usage: tools/gpurun_retry.sh <log-file> <gpurun args...>   -- retries while the pod answers "busy" (exit code 3)
log="$1"; shift
for attempt in $(seq 1 40); do
    /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
    rc=$?
    if [ $rc -ne 3 ]; then echo "gpurun rc=$rc after $attempt attempt(s)" >> "$log"; exit $rc; fi
    sleep 90
done
echo "gpurun still busy after 40 attempts" >> "$log"; exit 3
