#!/bin/bash
# every bench.py configuration on one GPU (short runs), one JSON line each -> gpurun_out/bench_<tag>_<config>.json
set -u
TAG=${1:-r02}
mkdir -p gpurun_out
for cfg in train infer aspp hrnet unet; do
  python bench.py --config $cfg --steps ${2:-20} --warmup 3 > gpurun_out/bench_${TAG}_${cfg}.json 2> gpurun_out/bench_${TAG}_${cfg}.err
  echo "$cfg rc=$?"; tail -3 gpurun_out/bench_${TAG}_${cfg}.err | cut -c1-300
  python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/bench_${TAG}_${cfg}.json').read().strip().splitlines()[-1])
    r = d.get('roofline', {})
    print('  %s: %.1f img/s, %.2f ms/step, e2e %.1f, launches %s, igemm frac %s exec %s, bn %s' % (
        '${cfg}', d['value'], d['ms_per_step'], d['e2e']['value'], d.get('gpu_launches'), r.get('frac'), r.get('frac_executed'),
        (d.get('roofline_bn') or {}).get('frac')))
except Exception as e:
    print('  parse failed:', e)
PY
done
