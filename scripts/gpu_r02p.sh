#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_round2.py tests/test_gpu_models.py tests/test_gpu_gradparity.py -q -m gpu -x > gpurun_out/pytest_r02p.log 2>&1
echo "rc=$?" | tee -a gpurun_out/pytest_r02p.log
for f in 0 1; do
B200SEG_FUSED_CE=$f python bench.py --steps 30 --warmup 5 > gpurun_out/bench_r02p_$f.json 2> gpurun_out/bench_r02p_$f.err
done
tail -4 gpurun_out/pytest_r02p.log
python - <<'PY'
import json
for f in (0,1):
    d=json.load(open('gpurun_out/bench_r02p_%d.json'%f))
    print(f, d['value'], d['ms_per_step'], d['e2e']['value'], d['config'].get('eager_images_per_sec'), d['config'].get('host_enqueue_ms_per_eager_step'))
PY
