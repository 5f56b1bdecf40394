#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_backbone.py tests/test_gpu_gradparity.py -q -x > gpurun_out/pytest_r02d.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_r02d.log
echo "== whole step"
for cfg in "1 1" "0 1" "1 0" "1 1"; do
  set -- $cfg
  B200SEG_IGEMM_DYNAMIC=$1 B200SEG_WGRAD_CLUSTER=$2 timeout 300 python bench.py --steps 40 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r02d_$1$2.json 2> gpurun_out/bench_r02d.err
  python - <<PY
import json
d = json.loads(open('gpurun_out/bench_r02d_$1$2.json').read().strip().splitlines()[-1])
print('dynamic=$1 cluster=$2: %.3f ms/step %.1f img/s e2e %.1f  igemm frac %.3f exec %.3f  bn frac %.3f (%.2f ms) host %.1f ms' % (
    d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['frac_executed'], d['roofline_bn']['frac'], d['roofline_bn']['ms_per_step'], d['config']['host_enqueue_ms_per_eager_step']))
PY
done
timeout 300 python scripts/bench_convs.py --set all --iters 10 > gpurun_out/convs_r02d.log 2>&1; grep -E "wgrad|aspp" gpurun_out/convs_r02d.log | head -40
