#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_jet_kernels.py -m gpu -x -q 2>&1 | tail -5
for n0 in 0 16 22 28 33; do GPE_DW_N0=$n0 timeout 300 python scripts/tc_kernel_times.py 4194304 3; done 2>&1 | tee gpurun_out/quick_times.log
