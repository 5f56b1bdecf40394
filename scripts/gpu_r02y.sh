#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_models.py tests/test_gpu_gradparity.py -x -q -m gpu > gpurun_out/pytest_r02y.log 2>&1
echo "rc=$?" | tee -a gpurun_out/pytest_r02y.log
tail -3 gpurun_out/pytest_r02y.log
for b in 0 1; do
B200SEG_WGRAD_BALANCE=$b python scripts/bench_convs.py --set head --iters 10 2>&1 | grep -E "wgrad" > gpurun_out/convs_bal$b.txt
done
paste -d'|' gpurun_out/convs_bal0.txt gpurun_out/convs_bal1.txt | cut -c1-46,80-125
bash scripts/gpu_ab_lib.sh "B200SEG_WGRAD_BALANCE=0 -" 40
