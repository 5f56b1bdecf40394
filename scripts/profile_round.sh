#!/bin/bash
# Profiling pass for one round (run on the GPU box through gpurun):
#   1. plain bench run (must exit 0 before anything runs under ncu)
#   2. ncu launch list (gpu__time_duration.sum, --clock-control none) of the same command
#   3. one `ncu --set full` capture of the tcgen05 conv kernels of one step
# Outputs land in gpurun_out/; summaries are copied into profiles/ by scripts/summarise_ncu.py here.
set -u
R=${1:-r01}
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${R}_short.log 2> gpurun_out/bench_${R}_short.err || { echo "bench failed"; tail -20 gpurun_out/bench_${R}_short.err; exit 1; }
cat gpurun_out/bench_${R}_short.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv \
    --log-file gpurun_out/launches_${R}.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list_${R}.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches_${R}.csv
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'igemm_kernel|wgrad_kernel' --launch-skip 0 -c 30 \
    -o gpurun_out/convs_${R} -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full_${R}.log 2>&1
echo "full capture rc=$?"; ls -la gpurun_out/
