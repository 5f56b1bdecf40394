#!/bin/bash
# A/B of the streaming BatchNorm passes: pixels per thread per trip (loads in flight) — scripts/bench_bandwidth.py rows
set -u
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_backbone.py -q -x -k "bn or batch or backbone or resnet or block" > gpurun_out/pytest_bn.log 2>&1; tail -3 gpurun_out/pytest_bn.log
for K in 1 2 4; do
  echo "== B200SEG_BN_K=$K"
  B200SEG_BN_K=$K python scripts/bench_bandwidth.py --iters 10 2>&1 | grep -E "^bn_|relu_bwd" | grep -E "16x128x128x256|16x32x32x1024|16x128x128x64"
done
