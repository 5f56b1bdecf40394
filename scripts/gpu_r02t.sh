#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_backbone.py tests/test_gpu_gradparity.py -x -q -m gpu > gpurun_out/pytest_r02t.log 2>&1
echo "rc=$?" | tee -a gpurun_out/pytest_r02t.log
python bench.py --steps 40 --warmup 5 > gpurun_out/bench_r02t.json 2> gpurun_out/bench_r02t.err
tail -3 gpurun_out/pytest_r02t.log
