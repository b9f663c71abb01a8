#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | head -40
timeout 400 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -c 600 gpurun_out/bench_n1.err
timeout 300 python bench.py --config 1 --steps 20 > gpurun_out/bench_cfg1.json 2> gpurun_out/bench_cfg1.err; echo "cfg1 rc=$?"; tail -c 600 gpurun_out/bench_cfg1.err
