#!/bin/bash
# usage: scripts/gpurun_retry.sh <timeout-seconds> '<command>'  — repeats the call while the pod answers "busy"/"transient"
T=$1; shift
for i in $(seq 1 20); do
    out=$(/usr/local/graft/bin/gpurun --timeout "$T" -- "$@" 2>&1); rc=$?
    echo "$out" | tail -60
    if [ $rc -eq 3 ] || echo "$out" | grep -q "status=transient"; then echo "[retry $i] rc=$rc"; sleep 90; continue; fi
    exit $rc
done
exit 3
