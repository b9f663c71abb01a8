#!/bin/bash
# One GPU visit: parity tests, bench line, ncu launch list, ncu --set full of the three tensor-path kernels.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
cat gpurun_out/bench_n1.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cat gpurun_out/bench_ref.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --points 4194304 > gpurun_out/ncu_launch.log 2>&1
python scripts/launch_summary.py gpurun_out/launches.csv
for k in jet_fwd_tc_kernel jet_rev_tc_kernel dw_tc_kernel; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -o gpurun_out/full_$k -f \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --points 1048576 > gpurun_out/ncu_full_$k.log 2>&1
done
ls -la gpurun_out
