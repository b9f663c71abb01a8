#!/bin/bash
mkdir -p gpurun_out
timeout 120 python scripts/fused_check.py 1048576 3 > gpurun_out/fused_check_1M.log 2>&1; echo "fused_check rc=$?"; tail -8 gpurun_out/fused_check_1M.log
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/bench_n1.json"))
    print({k: d[k] for k in ("value", "ms_per_step", "phase_ms", "clocks")}, d["roofline"].get("tensor_backward_ms"), d["e2e"]["value"])
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/bench_n1.err").read()[-2000:])
PY
