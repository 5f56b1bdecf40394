#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_gradparity.py -q -m gpu -k hrnet -s > gpurun_out/gradparity_r02n.log 2>&1
echo "hrnet gradparity rc=$?" | tee -a gpurun_out/gradparity_r02n.log
python scripts/profile_host.py > gpurun_out/host_profile_r02n.txt 2>&1
python bench.py --steps 30 --warmup 5 > gpurun_out/bench_r02n.json 2> gpurun_out/bench_r02n.err
grep -n "worst five\|passed\|failed" gpurun_out/gradparity_r02n.log | cut -c1-400
head -c 600 gpurun_out/bench_r02n.json; echo
head -70 gpurun_out/host_profile_r02n.txt
