#!/bin/bash
# developer helper: retry gpurun while the pod answers "busy" (exit code 3). Usage: tools/gpurun_retry.sh <log> <gpurun args...>
log=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc (attempt $i)" >> "$log"; exit $rc; fi
  sleep 90
done
echo "gpurun: still busy after 40 attempts" >> "$log"; exit 3
