#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_backbone.py -x -q -m gpu > gpurun_out/pytest_r02u.log 2>&1
echo "rc=$?" | tee -a gpurun_out/pytest_r02u.log
tail -3 gpurun_out/pytest_r02u.log
for k in 0 1; do
B200SEG_IGEMM_KPACK=$k python scripts/bench_convs.py --set backbone --iters 10 2>&1 | grep -E "^l1|^l2.conv|^l3.conv2" > gpurun_out/convs_kpack$k.txt
done
paste -d'|' gpurun_out/convs_kpack0.txt gpurun_out/convs_kpack1.txt | cut -c1-46,80-125
bash scripts/gpu_ab_lib.sh "B200SEG_IGEMM_KPACK=0 -" 40
