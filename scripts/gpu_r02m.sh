#!/bin/bash
# bandwidth-kernel rewrite: parity tests, HRNet mask-matched gradients, same-box A/B of the five kernels
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_backbone.py tests/test_gpu_models.py -q -m gpu -x > gpurun_out/pytest_r02m.log 2>&1
echo "kernels/backbone/models rc=$?" | tee -a gpurun_out/pytest_r02m.log
python -m pytest tests/test_gpu_gradparity.py -q -m gpu -k hrnet -s > gpurun_out/gradparity_r02m.log 2>&1
echo "hrnet gradparity rc=$?" | tee -a gpurun_out/gradparity_r02m.log
python scripts/bench_bw_kernels.py --lib pytorch_segmentation_b200/libb200seg_prev.so > gpurun_out/bw_prev.json 2> gpurun_out/bw_prev.err
python scripts/bench_bw_kernels.py > gpurun_out/bw_new.json 2> gpurun_out/bw_new.err
tail -3 gpurun_out/pytest_r02m.log; tail -30 gpurun_out/gradparity_r02m.log; cat gpurun_out/bw_prev.json gpurun_out/bw_new.json
