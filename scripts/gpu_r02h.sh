#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gradparity.py tests/test_gpu_trainstep.py tests/test_gpu_models.py -q -x > gpurun_out/pytest_r02h.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_r02h.log | cut -c1-200
for L in 3 1 2 4 3; do
  B200SEG_WGRAD_LANES=$L timeout 300 python bench.py --steps 40 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r02h_$L.json 2> gpurun_out/bench_r02h.err
  python - <<PY
import json
d = json.loads(open('gpurun_out/bench_r02h_$L.json').read().strip().splitlines()[-1])
print('lanes=$L: %.3f ms/step %.1f img/s e2e %.1f' % (d['ms_per_step'], d['value'], d['e2e']['value']))
PY
done
