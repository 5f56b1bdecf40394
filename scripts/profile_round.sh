#!/bin/bash
# Profiling pass for one round (run on the GPU box through gpurun):
#   1. plain bench run (must exit 0 before anything runs under ncu)
#   2. ncu launch list (gpu__time_duration.sum, --clock-control none) of the same command
#   3. `ncu --set full` captures of the dominant kernels of a steady-state step, summarised ON THE BOX with
#      scripts/summarise_full.py (gpurun_out/ may carry at most 64 MiB back: large .ncu-rep files are dropped after
#      they have been summarised; one small capture is kept for source-level reading)
# The profiled command is the eager step (--graph off): same kernels as the graph replay.
set -u
R=${1:-r01}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --graph off --no-cpu-baseline"
$CMD > gpurun_out/bench_${R}_short.log 2> gpurun_out/bench_${R}_short.err || { echo "bench failed"; tail -20 gpurun_out/bench_${R}_short.err; exit 1; }
cut -c1-300 gpurun_out/bench_${R}_short.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 14000 --csv \
    --log-file gpurun_out/launches_${R}.csv $CMD > gpurun_out/ncu_list_${R}.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches_${R}.csv
# steady state: skip the first three steps' worth of matching launches
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'igemm_kernel|wgrad_kernel' \
    --launch-skip 520 -c 60 -o /tmp/convs_${R} -f $CMD > gpurun_out/ncu_full_convs_${R}.log 2>&1
echo "conv capture rc=$?"
python scripts/summarise_full.py /tmp/convs_${R}.ncu-rep > gpurun_out/full_convs_${R}.md 2>> gpurun_out/ncu_full_convs_${R}.log
timeout 600 ncu --set full --clock-control none -k regex:'bn_bwd_reduce|bn_bwd_apply|bn_apply_kernel|grad_add_masked' \
    --launch-skip 520 -c 40 -o /tmp/bn_${R} -f $CMD > gpurun_out/ncu_full_bn_${R}.log 2>&1
echo "bn capture rc=$?"
python scripts/summarise_full.py /tmp/bn_${R}.ncu-rep > gpurun_out/full_bn_${R}.md 2>> gpurun_out/ncu_full_bn_${R}.log
# one small capture kept whole (source page): the ASPP 3x3 fprop and its neighbours
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'igemm_kernel' --launch-skip 575 -c 8 \
    -o gpurun_out/igemm_aspp_${R} -f $CMD > gpurun_out/ncu_full_small_${R}.log 2>&1
echo "small capture rc=$?"; du -sh gpurun_out; ls -la gpurun_out | tail -12
