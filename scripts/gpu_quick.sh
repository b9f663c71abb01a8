#!/bin/bash
timeout 300 python scripts/tc_kernel_times.py 4194304 3
timeout 300 python scripts/tc_check.py 2>&1 | grep -E "^d[12] "
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
