#!/bin/bash
# usage: scripts/gpu_retry.sh TIMEOUT [--gpus N] -- command ...   (retries while the pod answers "transient"/busy)
for attempt in $(seq 1 15); do
  /usr/local/graft/bin/gpurun --timeout "$@" > /tmp/gpurun_attempt.log 2>&1
  rc=$?
  if grep -q "status=transient\|status=busy" /tmp/gpurun_attempt.log || [ $rc -eq 3 ]; then
    echo "[gpu_retry] attempt $attempt: transient (rc $rc), retrying in 90 s" >&2
    sleep 90
    continue
  fi
  cat /tmp/gpurun_attempt.log
  exit $rc
done
cat /tmp/gpurun_attempt.log
exit 3
