#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_backbone.py tests/test_gpu_models.py -x -q -m gpu > gpurun_out/pytest_r02v.log 2>&1
echo "rc=$?" | tee -a gpurun_out/pytest_r02v.log
tail -3 gpurun_out/pytest_r02v.log
B200SEG_LIB_PATH=pytorch_segmentation_b200/libb200seg_prev.so python scripts/bench_convs.py --set backbone --iters 10 2>&1 | grep -E "fprop |dgrad" | grep -E "conv3|conv1|l3.conv2" > gpurun_out/convs_prev.txt
python scripts/bench_convs.py --set backbone --iters 10 2>&1 | grep -E "fprop |dgrad" | grep -E "conv3|conv1|l3.conv2" > gpurun_out/convs_new.txt
paste -d'|' gpurun_out/convs_prev.txt gpurun_out/convs_new.txt | cut -c1-46,80-125
bash scripts/gpu_ab_lib.sh "B200SEG_LIB_PATH=pytorch_segmentation_b200/libb200seg_prev.so -" 40
