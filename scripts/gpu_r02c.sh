#!/bin/bash
# wgrad cluster-reduce / direct-write and BN staged-load A/B: kernel tests, conv microbench, whole step
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_backbone.py -q -x > gpurun_out/pytest_r02c.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_r02c.log
for CL in 1 0; do
  echo "== wgrad microbench B200SEG_WGRAD_CLUSTER=$CL B200SEG_WGRAD_DIRECT=$CL"
  B200SEG_WGRAD_CLUSTER=$CL B200SEG_WGRAD_DIRECT=$CL timeout 300 python scripts/bench_convs.py --set all --iters 10 2>&1 | grep -E "wgrad" | awk '{printf "%-16s %8s us %8s TF\n", $1, $3, $4}'
done
echo "== whole step"
for cfg in "1 1 4 2 2 32" "0 0 4 2 2 32" "1 1 1 1 1 32" "1 1 4 2 2 256"; do
  set -- $cfg
  B200SEG_WGRAD_CLUSTER=$1 B200SEG_WGRAD_DIRECT=$2 B200SEG_BN_K_APPLY=$3 B200SEG_BN_K_RED=$4 B200SEG_BN_K_BWD=$5 B200SEG_BNRED_GPB=$6 \
    timeout 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r02c_$1$2$3$4$5_$6.json 2> gpurun_out/bench_r02c.err
  python - <<PY
import json
d = json.loads(open('gpurun_out/bench_r02c_$1$2$3$4$5_$6.json').read().strip().splitlines()[-1])
print('cluster=$1 direct=$2 bnK=$3/$4/$5 gpb=$6: %.3f ms/step %.1f img/s  igemm frac %.3f exec %.3f  bn frac %.3f (%.2f ms)' % (
    d['ms_per_step'], d['value'], d['roofline']['frac'], d['roofline']['frac_executed'], d['roofline_bn']['frac'], d['roofline_bn']['ms_per_step']))
PY
done
