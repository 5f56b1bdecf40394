#!/bin/bash
# gpurun with retries while the pod answers busy (exit 3): scripts/gpurun_retry.sh <timeout> <out-file> <command...>
T=$1; OUT=$2; shift 2
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $T -- "$@" > "$OUT" 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" "$OUT"; then exit $rc; fi
  sleep 45
done
exit 3
