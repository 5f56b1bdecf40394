#!/bin/bash
# operand-role swap in igemm_kernel: parity with the swap forced wherever legal, then where it pays; timing A/B
mkdir -p gpurun_out
B200SEG_IGEMM_SWAP=2 timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_backbone.py -x -q -m gpu > gpurun_out/pytest_swap2.log 2>&1
echo "forced rc=$?"; tail -4 gpurun_out/pytest_swap2.log | cut -c1-300
timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "conv_fprop_dgrad_wgrad" > gpurun_out/pytest_swap1.log 2>&1
echo "auto rc=$?"; tail -3 gpurun_out/pytest_swap1.log | cut -c1-300
for b in 0 1; do
B200SEG_IGEMM_SWAP=$b timeout 300 python scripts/bench_convs.py --set backbone --iters 10 2>&1 | grep -E "^l1" | grep -E "fprop|dgrad" > gpurun_out/convs_swap$b.txt
done
paste -d'|' gpurun_out/convs_swap0.txt gpurun_out/convs_swap1.txt | cut -c1-46,80-125
timeout 600 bash scripts/gpu_ab_lib.sh "B200SEG_IGEMM_SWAP=0 -" 40
