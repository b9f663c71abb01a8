#!/bin/bash
# one `ncu --set full` capture of the three tensor-path backward kernels at the chunk size (1048576 points = one backward chunk), after the same
# command has exited 0 without ncu
mkdir -p gpurun_out
timeout 300 python scripts/tc_kernel_times.py 1048576 1 > gpurun_out/r2g_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/r2g_plain.log; exit 1; }
tail -1 gpurun_out/r2g_plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'jet_fwd2_tc_kernel|jet_rev2_tc_kernel|dw_tc_kernel' -s 9 -c 3 \
    -o gpurun_out/r2g_bwd3 -f python scripts/tc_kernel_times.py 1048576 1 > gpurun_out/r2g_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r2g_ncu.log; ls -la gpurun_out/r2g_bwd3.ncu-rep
