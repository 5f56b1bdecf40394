#!/bin/bash
set -u
mkdir -p gpurun_out
for S in 0 1 0; do
  B200SEG_SKIP_ALLREDUCE=$S timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29580 + S)) \
     bench.py --gpus 2 --steps 40 --warmup 3 > gpurun_out/bench_n2_skip$S.json 2> gpurun_out/bench_n2_skip.err
  python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/bench_n2_skip$S.json').read().strip().splitlines()[-1])
    print('skip_allreduce=$S: %.3f ms/step %.1f img/s sync %s' % (d['ms_per_step'], d['value'], d.get('param_sync_max_abs_diff')))
except Exception as e:
    print('failed', e); print(open('gpurun_out/bench_n2_skip.err').read()[-800:])
PY
done
timeout 300 python bench.py --steps 40 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('N=1 same box: %.3f ms/step' % d['ms_per_step'])"
