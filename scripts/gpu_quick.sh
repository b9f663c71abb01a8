#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6; exit 0
python bench.py --steps 3 --no-cpu-baseline > gpurun_out/bench_stash.json 2> gpurun_out/bench_stash.err; echo rc=$?; tail -3 gpurun_out/bench_stash.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_stash.json'))
print(d['value'], d['ms_per_step'], d['phase_ms'], d['e2e']['value'], d['clocks'], d['roofline']['tensor_backward_ms'])
PY
GPE_STASH_MAX_GB=0 python bench.py --steps 3 --no-cpu-baseline | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('recompute:', d['value'], d['ms_per_step'], d['phase_ms'])"
nvidia-smi --query-gpu=memory.used,memory.total --format=csv
