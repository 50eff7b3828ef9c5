#!/usr/bin/env bash
# tools/gpurun_retry.sh <timeout_s> <command...>: gpurun, retried while the pod answers busy / transient (nothing charged)
T=$1; shift
for i in $(seq 1 40); do
  out=$(/usr/local/graft/bin/gpurun --timeout "$T" -- "$@" 2>&1)
  echo "$out" | tail -150
  if echo "$out" | grep -q "status=transient\|rc=3\|no box\|busy"; then
    if echo "$out" | grep -q "status=ok"; then exit 0; fi
    echo "[retry $i] waiting 90 s"; sleep 90; continue
  fi
  exit 0
done
