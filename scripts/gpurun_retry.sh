#!/bin/bash
# usage: scripts/gpurun_retry.sh <timeout_s> <command...>   -- retries while the pod answers "transient" / busy
t=$1; shift
for try in $(seq 1 40); do
    out=$(/usr/local/graft/bin/gpurun --timeout $t -- "$@" 2>&1)
    if echo "$out" | grep -q "status=transient\|exit code 3\|rc=3 "; then
        echo "[retry $try] pod busy; sleeping" >&2
        sleep 150
        continue
    fi
    echo "$out"
    exit 0
done
echo "gave up after 40 tries"
