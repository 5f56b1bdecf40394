#!/bin/bash
# ncu launch list (gpu__time_duration.sum, --clock-control none) of the eager training step, after the plain run exits 0
set -u
R=${1:-r02}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --graph off --no-cpu-baseline"
$CMD > gpurun_out/bench_${R}_short.log 2> gpurun_out/bench_${R}_short.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 16000 --csv \
    --log-file gpurun_out/launches_${R}.csv $CMD > gpurun_out/ncu_list_${R}.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches_${R}.csv
python scripts/summarise_ncu.py gpurun_out/launches_${R}.csv --steps 13 > gpurun_out/launches_${R}.md; head -40 gpurun_out/launches_${R}.md
gzip -f gpurun_out/launches_${R}.csv
