#!/bin/bash
# the allocation-free entry point: alone in a fresh process, then at the end of the whole GPU suite
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -q -m gpu -rs > gpurun_out/pytest_noalloc_alone.log 2>&1
echo "alone rc=$?" | tee -a gpurun_out/pytest_noalloc_alone.log
python -m pytest tests -x -q -m gpu -rs > gpurun_out/pytest_gpu_all.log 2>&1
echo "all rc=$?" | tee -a gpurun_out/pytest_gpu_all.log
tail -8 gpurun_out/pytest_noalloc_alone.log; tail -8 gpurun_out/pytest_gpu_all.log
