#!/bin/bash
# every named BASELINE config at 1 GPU with the final code: one JSON line each
mkdir -p gpurun_out
for c in 4 2 3 5 1; do
  timeout 500 python bench.py --config $c > gpurun_out/r2g_cfg${c}_n1.json 2> gpurun_out/r2g_cfg${c}_n1.err; echo "cfg $c rc=$?"
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/r2g_cfg${c}_n1.json"))
    print({k: d.get(k) for k in ("value", "ms_per_step")}, d["e2e"]["value"], d["roofline"].get("frac"), d["clocks"], (d.get("graph_replay") or {}).get("ms_per_step"))
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/r2g_cfg${c}_n1.err").read()[-800:])
PY
done
