#!/bin/bash
# same-box A/B of an environment switch at step level: scripts/ab_env.sh VAR v1 v2 [v1 v2 ...]
VAR=$1; shift
for v in "$@"; do
  echo "$VAR=$v $(env $VAR=$v python bench.py --steps 60 --warmup 8 --no-cpu-baseline 2>/dev/null | tail -1 | grep -o '"ms_per_step": [0-9.]*' | head -1)"
done
