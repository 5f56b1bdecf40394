#!/bin/bash
# Round-2 profiling pass (final kernel state): launch list of one eager training step, ncu --set full of the head convs
# (the ones the metric names) and of every igemm launch of one whole step (DRAM traffic per launch -> igemm_traffic.json)
set -u
R=${1:-r02}
mkdir -p gpurun_out
bash scripts/gpu_launchlist.sh $R
python scripts/profile_head.py --tags gpurun_out/head_tags_${R}.json > gpurun_out/head_plain_${R}.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'igemm_kernel|wgrad_kernel' \
    -o /tmp/head_${R} -f python scripts/profile_head.py > gpurun_out/ncu_head_${R}.log 2>&1
echo "head capture rc=$?"
python scripts/summarise_full.py /tmp/head_${R}.ncu-rep --tags gpurun_out/head_tags_${R}.json > gpurun_out/full_head_${R}.md 2>> gpurun_out/ncu_head_${R}.log
cat gpurun_out/full_head_${R}.md | head -30
CMD="python bench.py --steps 2 --warmup 3 --graph off --no-cpu-baseline"
timeout 1200 ncu --set full --clock-control none -k regex:'igemm_kernel' --launch-skip 1116 -c 124 -o /tmp/igemm_${R} -f $CMD > gpurun_out/ncu_igemm_${R}.log 2>&1
echo "igemm capture rc=$?"
python scripts/summarise_full.py /tmp/igemm_${R}.ncu-rep > gpurun_out/full_igemm_${R}.md 2>> gpurun_out/ncu_igemm_${R}.log
tail -5 gpurun_out/full_igemm_${R}.md
timeout 900 ncu --set full --clock-control none -k regex:'bn_bwd_reduce|bn_bwd_apply|bn_apply_kernel|grad_add_masked|grad_join_reduce' \
    --launch-skip 1800 -c 60 -o /tmp/bn_${R} -f $CMD > gpurun_out/ncu_bn_${R}.log 2>&1
echo "bn capture rc=$?"
python scripts/summarise_full.py /tmp/bn_${R}.ncu-rep > gpurun_out/full_bn_${R}.md 2>> gpurun_out/ncu_bn_${R}.log
tail -5 gpurun_out/full_bn_${R}.md
