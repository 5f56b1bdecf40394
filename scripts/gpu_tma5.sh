#!/bin/bash
# 5-D TMA operand loads in wgrad_kernel: parity, per-shape timing with the switch off / on, whole-step A/B
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_backbone.py tests/test_gpu_models.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/pytest_tma5.log 2>&1
echo "rc=$?" | tee -a gpurun_out/pytest_tma5.log
tail -3 gpurun_out/pytest_tma5.log
for b in 0 1; do
B200SEG_WGRAD_TMA5=$b python scripts/bench_convs.py --set all --iters 10 2>&1 | grep -E "wgrad" > gpurun_out/convs_tma5_$b.txt
done
paste -d'|' gpurun_out/convs_tma5_0.txt gpurun_out/convs_tma5_1.txt | cut -c1-46,80-125
bash scripts/gpu_ab_lib.sh "B200SEG_WGRAD_TMA5=0 -" 40
