#!/bin/bash
set -u
mkdir -p gpurun_out
for cfg in "1 0" "0 1" "0 0" "1 0" "0 1"; do
  set -- $cfg
  B200SEG_IGEMM_DYNAMIC=$1 B200SEG_IGEMM_NFAST=$2 timeout 300 python bench.py --steps 40 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r02g_$1$2.json 2> gpurun_out/bench_r02g.err
  python - <<PY
import json
d = json.loads(open('gpurun_out/bench_r02g_$1$2.json').read().strip().splitlines()[-1])
print('dynamic=$1 nfast=$2: %.3f ms/step %.1f img/s  igemm frac %.3f exec %.3f' % (d['ms_per_step'], d['value'], d['roofline']['frac'], d['roofline']['frac_executed']))
PY
done
