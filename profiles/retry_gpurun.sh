#!/bin/bash
# usage: retry_gpurun.sh OUTFILE gpurun-args...   (retries while the pod answers transient/busy)
out=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > $out 2>&1
  if grep -q "status=transient" $out; then sleep 150; else break; fi
done
