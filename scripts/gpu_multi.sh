#!/bin/bash
# multi-GPU visit: sharded == single-GPU over NCCL, then the bench line (weak value + strong record) at N ranks
N=${1:-2}
mkdir -p gpurun_out
P=$((29700 + N))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P scripts/nccl_equivalence.py > gpurun_out/nccl_equivalence_n$N.json 2> gpurun_out/nccl_equivalence_n$N.err; echo "equiv rc=$?"; cat gpurun_out/nccl_equivalence_n$N.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((P+10)) bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_n$N.json"))
    print({k: d[k] for k in ("value", "ms_per_step", "n_gpus", "strong")}, d["e2e"]["value"])
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/bench_n$N.err").read()[-1500:])
PY
