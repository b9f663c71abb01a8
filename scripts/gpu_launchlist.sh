#!/bin/bash
# launch list of the bench command (4M points so that 700 launches cover warm-up + 2 steps), after the same command ran without ncu
mkdir -p gpurun_out
timeout 400 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --points 4194304 > gpurun_out/r2g_bench_4M.json 2> gpurun_out/r2g_bench_4M.err || { echo plain failed; tail -5 gpurun_out/r2g_bench_4M.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r2g_tc_launches_bench_steps2_4M.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --points 4194304 > gpurun_out/r2g_launch_ncu.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/r2g_tc_launches_bench_steps2_4M.csv
