#!/bin/bash
# 2-GPU sweep of NCCL's CTA budget (the all-reduce runs under backward and takes SMs from the persistent conv kernels)
set -u
mkdir -p gpurun_out
for C in 0 4 8 16; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29520 + C)) \
     bench.py --gpus 2 --steps 30 --warmup 3 --nccl-ctas $C > gpurun_out/bench_n2_ctas$C.json 2> gpurun_out/bench_n2_ctas$C.err
  python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/bench_n2_ctas$C.json').read().strip().splitlines()[-1])
    print('nccl_ctas=$C: %.3f ms/step %.1f img/s e2e %.1f sync %s' % (d['ms_per_step'], d['value'], d['e2e']['value'], d.get('param_sync_max_abs_diff')))
except Exception as e:
    print('nccl_ctas=$C failed', e)
PY
done
