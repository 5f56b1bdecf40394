#!/bin/bash
# same-box A/B of one environment switch: scripts/gpu_ab_env.sh VAR "v1 v2 ..." [steps]
set -u
VAR=$1; VALS=$2; STEPS=${3:-40}
mkdir -p gpurun_out
for rep in 1 2; do
for v in $VALS; do
  env $VAR=$v timeout 300 python bench.py --steps $STEPS --warmup 3 --no-cpu-baseline > gpurun_out/bench_ab_${VAR}_$v.json 2> gpurun_out/bench_ab.err
  python - <<PY
import json
d = json.loads(open('gpurun_out/bench_ab_${VAR}_$v.json').read().strip().splitlines()[-1])
print('$VAR=$v: %.3f ms/step %.1f img/s  igemm frac %.3f exec %.3f bn %.2f ms' % (d['ms_per_step'], d['value'], d['roofline']['frac'], d['roofline']['frac_executed'], d['roofline_bn']['ms_per_step']))
PY
done
done
