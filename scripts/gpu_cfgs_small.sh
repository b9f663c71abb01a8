#!/bin/bash
# the host-bound configs (eager launch trains): run twice, keep both (the box's host cores are shared; see profiles/README.md)
mkdir -p gpurun_out
for c in 2 3 1; do for r in a b; do
  timeout 500 python bench.py --config $c > gpurun_out/r2g_cfg${c}_n1_$r.json 2> gpurun_out/r2g_cfg${c}_n1_$r.err; echo "cfg $c run $r rc=$?"
  python - <<PY
import json
d = json.load(open("gpurun_out/r2g_cfg${c}_n1_$r.json"))
print({k: d.get(k) for k in ("value", "ms_per_step")}, d["e2e"]["value"], d["phase_ms"], (d.get("graph_replay") or {}).get("ms_per_step"))
PY
done; done
