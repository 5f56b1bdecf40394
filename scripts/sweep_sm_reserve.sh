#!/bin/bash
# N-GPU sweeps of the data-parallel knobs (run through gpurun --gpus N): SMs left to NCCL, NCCL CTA cap, bucket size.
N=${1:-4}
P=29530
run() {
  P=$((P+1))
  echo "$*: $(timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline "$@" 2>>gpurun_out/sweep_dp.err | tail -1 | cut -c60-190)"
}
if [ "${2:-buckets}" = "reserve" ]; then
  for cfg in "0 0" "8 8" "16 16" "0 8" "16 0" "24 24"; do set -- $cfg; run --sm-reserve $1 --nccl-ctas $2; done
else
  for mb in 25 8 64 200; do run --bucket-mb $mb; done
fi
