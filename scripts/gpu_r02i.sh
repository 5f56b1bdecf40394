#!/bin/bash
set -u
mkdir -p gpurun_out
for C in 8 4 2 8; do
  B200SEG_WGRAD_MAXCS=$C timeout 300 python bench.py --steps 40 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r02i_$C.json 2> gpurun_out/bench_r02i.err
  python - <<PY
import json
d = json.loads(open('gpurun_out/bench_r02i_$C.json').read().strip().splitlines()[-1])
print('max cluster=$C: %.3f ms/step %.1f img/s' % (d['ms_per_step'], d['value']))
PY
done
