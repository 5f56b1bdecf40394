#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_r02e.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_r02e.log
echo "== whole step"
for cfg in "1" "0" "1"; do
  B200SEG_JOIN_REDUCE=$cfg timeout 300 python bench.py --steps 40 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r02e_$cfg.json 2> gpurun_out/bench_r02e.err
  python - <<PY
import json
d = json.loads(open('gpurun_out/bench_r02e_$cfg.json').read().strip().splitlines()[-1])
print('join_reduce=$cfg: %.3f ms/step %.1f img/s e2e %.1f  igemm frac %.3f exec %.3f  bn frac %.3f (%.2f ms) launches %s' % (
    d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['frac_executed'], d['roofline_bn']['frac'], d['roofline_bn']['ms_per_step'], d['gpu_launches']))
PY
done
