#!/bin/bash
# gpurun with retries while the pod answers "transient" (nothing charged): tools/gpurun_retry.sh TIMEOUT 'command' [LOG]
T=$1; CMD=$2; LOG=${3:-gpurun_out/retry.log}
mkdir -p gpurun_out
for i in $(seq 1 12); do
  /usr/local/graft/bin/gpurun --timeout "$T" -- "$CMD" > "$LOG" 2>&1
  rc=$?
  if ! grep -q "status=transient" "$LOG"; then exit $rc; fi
  sleep 240
done
exit 3
