#!/bin/bash
# which of the round-2 kernel changes makes the gradient-parity suite flaky: repeat it under each toggle
set -u
run() {
  local tag="$1"; shift
  local fails=0
  for i in 1 2 3 4; do
    env "$@" python -m pytest tests/test_gpu_gradparity.py -q -x 2>&1 | grep -E "passed|failed" | tail -1 > /tmp/r.txt
    if grep -q failed /tmp/r.txt; then fails=$((fails+1)); fi
  done
  echo "$tag: $fails / 4 runs failed"
}
run "all new (default)" X=1
run "dynamic scheduler off" B200SEG_IGEMM_DYNAMIC=0
run "wgrad cluster+direct off" B200SEG_WGRAD_CLUSTER=0 B200SEG_WGRAD_DIRECT=0
run "BN K=1" B200SEG_BN_K=1
run "join fusion off" B200SEG_JOIN_REDUCE=0
run "everything off" B200SEG_IGEMM_DYNAMIC=0 B200SEG_WGRAD_CLUSTER=0 B200SEG_WGRAD_DIRECT=0 B200SEG_BN_K=1 B200SEG_JOIN_REDUCE=0 B200SEG_BNRED_GPB=256
