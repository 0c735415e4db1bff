#!/bin/bash
# usage: gpurun_retry.sh <out file> <gpurun args...>   retries while the pod answers "busy" (rc 3), max ~40 min
OUT=$1; shift
for i in $(seq 1 16); do
  /usr/local/graft/bin/gpurun "$@" > "$OUT" 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" "$OUT"; then echo "rc=$rc" >> "$OUT"; exit $rc; fi
  sleep 120
done
echo "gave up" >> "$OUT"
