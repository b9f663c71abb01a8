#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench.py --config 1 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2g_cfg1_plain.json 2> gpurun_out/r2g_cfg1_plain.err || { echo plain failed; tail -5 gpurun_out/r2g_cfg1_plain.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2g_cfg1_launches.csv \
    python bench.py --config 1 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2g_cfg1_ncu.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/r2g_cfg1_launches.csv
