#!/bin/bash
set -u
mkdir -p gpurun_out
for cfg in "64 4" "64 0" "64 1" "25 4" "64 12"; do
  set -- $cfg
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29540 + $2)) \
     bench.py --gpus 2 --steps 30 --warmup 3 --bucket-mb $1 --last-bucket-mb $2 > gpurun_out/bench_n2_b$1_l$2.json 2> gpurun_out/bench_n2_b.err
  python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/bench_n2_b$1_l$2.json').read().strip().splitlines()[-1])
    print('bucket=$1 last=$2: %.3f ms/step %.1f img/s e2e %.1f sync %s' % (d['ms_per_step'], d['value'], d['e2e']['value'], d.get('param_sync_max_abs_diff')))
except Exception as e:
    print('failed', e)
PY
done
