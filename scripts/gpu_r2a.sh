#!/bin/bash
# round-2 GPU visit A: parity suite + precision check + per-kernel times + bench line (every step under its own timeout)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python scripts/tc_check.py > gpurun_out/tc_check.log 2>&1; echo "tc_check rc=$?"; tail -25 gpurun_out/tc_check.log
timeout 200 python scripts/tc_kernel_times.py 4194304 3 > gpurun_out/tc_times.log 2>&1; echo "times rc=$?"; cat gpurun_out/tc_times.log
timeout 400 python bench.py --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
cat gpurun_out/bench_n1.json
