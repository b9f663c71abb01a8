#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_jet_kernels.py -m gpu -x -q 2>&1 | tail -3
for n0 in 0 21 25 29 33 37; do GPE_DW_N0=$n0 python scripts/tc_kernel_times.py 4194304 3; done 2>&1 | tee gpurun_out/quick_times.log
python scripts/tc_bign.py 2>&1 | tee gpurun_out/bign.log
