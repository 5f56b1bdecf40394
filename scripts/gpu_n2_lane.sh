#!/bin/bash
set -u
mkdir -p gpurun_out
for L in 1 0 1; do
  B200SEG_SINK_ON_LANE=$L timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29560 + L)) \
     bench.py --gpus 2 --steps 40 --warmup 3 > gpurun_out/bench_n2_lane$L.json 2> gpurun_out/bench_n2_lane.err
  python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/bench_n2_lane$L.json').read().strip().splitlines()[-1])
    print('sink_on_lane=$L: %.3f ms/step %.1f img/s e2e %.1f sync %s exposed %s' % (d['ms_per_step'], d['value'], d['e2e']['value'], d.get('param_sync_max_abs_diff'), d.get('allreduce_exposed_ms_per_step')))
except Exception as e:
    print('failed', e); print(open('gpurun_out/bench_n2_lane.err').read()[-1500:])
PY
done
