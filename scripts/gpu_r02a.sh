#!/bin/bash
# round-2 first GPU call: baseline state of tests / bench, and the ncu --set full capture of the head convs
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_r02a.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_r02a.log
python bench.py --steps 20 --warmup 3 --detail gpurun_out/detail_r02a.json > gpurun_out/bench_r02a.log 2> gpurun_out/bench_r02a.err; echo "bench rc=$?"
cut -c1-400 gpurun_out/bench_r02a.log
python scripts/profile_head.py --tags gpurun_out/head_tags_r02a.json > gpurun_out/head_plain_r02a.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'igemm_kernel|wgrad_kernel' \
    -o /tmp/head_r02a -f python scripts/profile_head.py > gpurun_out/ncu_head_r02a.log 2>&1
echo "head capture rc=$?"; tail -30 gpurun_out/head_plain_r02a.log
python scripts/summarise_full.py /tmp/head_r02a.ncu-rep --tags gpurun_out/head_tags_r02a.json > gpurun_out/full_head_r02a.md 2>> gpurun_out/ncu_head_r02a.log
ls -la /tmp/head_r02a.ncu-rep; cat gpurun_out/full_head_r02a.md | head -50
