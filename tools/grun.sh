#!/bin/bash
# Retry wrapper around gpurun (the pod answers "transient" while it drains other requests): tools/grun.sh <timeout_s> [--gpus N] -- '<command>'
T=$1; shift
for i in 1 2 3 4 5 6 7 8 9 10 11 12; do
    out=$(timeout $((T + 2400)) gpurun --timeout $T "$@" 2>&1)
    if echo "$out" | grep -q "status=transient\|status=busy\|status=refused"; then echo "[grun] attempt $i: busy, retrying in 60 s"; sleep 60; continue; fi
    echo "$out"; exit 0
done
echo "$out"; exit 3
