#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_jet_kernels.py -m gpu -x -q 2>&1 | tail -8
for n1 in 0 46 53 58; do GPE_DW_N1=$n1 timeout 300 python scripts/tc_kernel_times.py 4194304 3; done 2>&1 | tee gpurun_out/quick_times.log
timeout 600 python scripts/tc_bign.py 2>&1 | tee gpurun_out/bign.log
