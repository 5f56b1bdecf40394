set -u
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r01b_short.log 2> gpurun_out/bench_r01b_short.err || { echo "bench failed"; tail -20 gpurun_out/bench_r01b_short.err; exit 1; }
cut -c1-300 gpurun_out/bench_r01b_short.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv \
    --log-file gpurun_out/launches_r01b.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list_r01b.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches_r01b.csv
