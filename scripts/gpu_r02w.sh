#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_backbone.py tests/test_gpu_models.py -x -q -m gpu > gpurun_out/pytest_r02w.log 2>&1
echo "rc=$?" | tee -a gpurun_out/pytest_r02w.log
tail -3 gpurun_out/pytest_r02w.log
bash scripts/gpu_ab_lib.sh "B200SEG_LIB_PATH=pytorch_segmentation_b200/libb200seg_prev.so -" 40
