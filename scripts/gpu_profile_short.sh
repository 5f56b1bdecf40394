#!/bin/bash
# launch list of one eager training step + ncu --set full of the head convs only (the igemm / BN captures of
# scripts/gpu_profile_r02.sh are unchanged when only wgrad_kernel changed)
set -u
R=${1:-r02g}
mkdir -p gpurun_out
bash scripts/gpu_launchlist.sh $R
python scripts/profile_head.py --tags gpurun_out/head_tags_${R}.json > gpurun_out/head_plain_${R}.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'igemm_kernel|wgrad_kernel' \
    -o /tmp/head_${R} -f python scripts/profile_head.py > gpurun_out/ncu_head_${R}.log 2>&1
echo "head capture rc=$?"
python scripts/summarise_full.py /tmp/head_${R}.ncu-rep --tags gpurun_out/head_tags_${R}.json > gpurun_out/full_head_${R}.md 2>> gpurun_out/ncu_head_${R}.log
cat gpurun_out/full_head_${R}.md | head -30
