#!/bin/bash
# Profiling pass for one round (run on the GPU box through gpurun):
#   1. plain bench run (must exit 0 before anything runs under ncu)
#   2. ncu launch list (gpu__time_duration.sum, --clock-control none) of the same command
#   3. one `ncu --set full` capture of the dominant kernels of one steady-state step
# The profiled command is the eager step (--graph off): same kernels as the graph replay, one process, one stream
# plus the weight-gradient lane.  Outputs land in gpurun_out/; summaries are made here by scripts/summarise_ncu.py and
# scripts/summarise_full.py and committed under profiles/.
set -u
R=${1:-r01}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --graph off --no-cpu-baseline"
$CMD > gpurun_out/bench_${R}_short.log 2> gpurun_out/bench_${R}_short.err || { echo "bench failed"; tail -20 gpurun_out/bench_${R}_short.err; exit 1; }
cut -c1-600 gpurun_out/bench_${R}_short.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 14000 --csv \
    --log-file gpurun_out/launches_${R}.csv $CMD > gpurun_out/ncu_list_${R}.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches_${R}.csv
timeout 900 ncu --set full --clock-control none --import-source on \
    -k regex:'igemm_kernel|wgrad_kernel|bn_bwd_reduce|bn_bwd_apply|bn_apply_kernel' --launch-skip 1200 -c 120 \
    -o gpurun_out/step_${R} -f $CMD > gpurun_out/ncu_full_${R}.log 2>&1
echo "full capture rc=$?"; ls -la gpurun_out/ | tail -5
