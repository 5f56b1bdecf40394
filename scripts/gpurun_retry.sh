#!/bin/bash
# gpurun with retries while the pod answers busy (exit 3): scripts/gpurun_retry.sh <timeout> <out-file> <command...>
# GPUS=<n> in the environment asks for n GPUs of one box (charged n x)
T=$1; OUT=$2; shift 2
G=""
if [ -n "${GPUS:-}" ]; then G="--gpus $GPUS"; fi
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun $G --timeout $T -- "$@" > "$OUT" 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" "$OUT"; then exit $rc; fi
  sleep 45
done
exit 3
