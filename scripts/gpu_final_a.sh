#!/bin/bash
# final validation: whole GPU suite, smoke, every bench configuration, the reference arm
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_final.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/pytest_final.log
tail -3 gpurun_out/pytest_final.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1
echo "smoke rc=$?"; tail -4 gpurun_out/smoke_final.log | cut -c1-300
bash scripts/gpu_bench_all.sh r02f 30
python bench.py --steps 40 --warmup 5 > gpurun_out/bench_r02f_train_full.json 2> gpurun_out/bench_r02f_train_full.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r02f_reference.json 2> gpurun_out/bench_r02f_reference.err
head -c 400 gpurun_out/bench_r02f_reference.json; echo
