#!/bin/bash
mkdir -p gpurun_out
for il in 0 1; do
  echo "=== GPE_TC_INTERLEAVE=$il"
  GPE_TC_INTERLEAVE=$il timeout 300 python scripts/tc_kernel_times.py 4194304 3
  GPE_TC_INTERLEAVE=$il timeout 300 python scripts/tc_check.py 2>&1 | grep -E "^d[12] "
  GPE_TC_INTERLEAVE=$il timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
done 2>&1 | tee gpurun_out/interleave.log
