#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_round2.py -q -m gpu -x > gpurun_out/pytest_r02o.log 2>&1
echo "round2 rc=$?" | tee -a gpurun_out/pytest_r02o.log
python scripts/profile_host.py --top 25 > gpurun_out/host_profile_r02o.txt 2>&1
python bench.py --steps 30 --warmup 5 > gpurun_out/bench_r02o.json 2> gpurun_out/bench_r02o.err
tail -5 gpurun_out/pytest_r02o.log
head -3 gpurun_out/host_profile_r02o.txt | cut -c1-300
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_r02o.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['config'].get('eager_images_per_sec'), d['config'].get('host_enqueue_ms_per_eager_step'))
PY
