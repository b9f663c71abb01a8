#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_jet_kernels.py -m gpu -x -q 2>&1 | tail -3
for pf in 0 2 4 6 8 12; do GPE_DW_PF=$pf timeout 300 python scripts/tc_kernel_times.py 4194304 3; done 2>&1 | tee gpurun_out/quick_times.log
