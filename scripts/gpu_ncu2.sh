#!/bin/bash
mkdir -p gpurun_out
for k in jet_rev_tc_kernel dw_tc_kernel; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -o gpurun_out/full2_$k -f \
    python scripts/tc_kernel_times.py 1048576 1 > gpurun_out/ncu_full2_$k.log 2>&1
done
ls -la gpurun_out | tail -5
