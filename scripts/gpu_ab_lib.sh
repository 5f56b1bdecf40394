#!/bin/bash
# same-box A/B of bench.py under a list of "ENV=VAL[,ENV=VAL...]" settings ("-" = defaults), two rounds:
#   scripts/gpu_ab_lib.sh "- B200SEG_LIB_PATH=pytorch_segmentation_b200/libb200seg_prev.so B200SEG_BNAPPLY_GPB=32" [steps]
set -u
SETS=$1; STEPS=${2:-40}
mkdir -p gpurun_out
for rep in 1 2; do
for s in $SETS; do
  if [ "$s" = "-" ]; then E=""; else E=$(echo $s | tr ',' ' '); fi
  env $E timeout 300 python bench.py --steps $STEPS --warmup 3 --no-cpu-baseline > gpurun_out/bench_ab_cur.json 2> gpurun_out/bench_ab.err
  python - "$s" <<'PY'
import json, sys
d = json.loads(open('gpurun_out/bench_ab_cur.json').read().strip().splitlines()[-1])
print('%-70s %.3f ms/step %.1f img/s  igemm frac %.3f  bn %.2f ms (%.0f GB/s)' % (sys.argv[1], d['ms_per_step'], d['value'], d['roofline']['frac'], d['roofline_bn']['ms_per_step'], d['roofline_bn']['achieved']))
PY
done
done
