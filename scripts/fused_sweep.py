"""Sweep the fused backward's split (reverse-sweep clusters) and ring size at the config-4 shape.  usage: fused_sweep.py [n] [reps]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import torch
from gpe_b200 import ops
from test_jet_kernels import make_net
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 22
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
d, W, L, jet = 2, 128, 4, 3
Ws, bs = make_net(d, W, L, seed=1, act=0)
Ws = [w.cuda() for w in Ws]; bs = [b.cuda() for b in bs]
x = torch.rand(n, d, device="cuda") * 4 - 2
ub = torch.randn(n, device="cuda"); gb = torch.randn(n, d, device="cuda"); hb = torch.randn(n, 1, device="cuda")
u, gg, h, stash = ops.mlp_jet_fwd_stash(x, Ws, bs, 0, jet)
def timed():
    for _ in range(2):
        ops.mlp_jet_bwd_stashed(x, Ws, bs, 0, jet, ub, gb, hb, stash)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(reps):
        ops.mlp_jet_bwd_stashed(x, Ws, bs, 0, jet, ub, gb, hb, stash)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps
os.environ["GPE_BWD_FUSED"] = "0"
print(json.dumps({"n": n, "chunked_ms": timed()}), flush=True)
os.environ["GPE_BWD_FUSED"] = "1"
combos = [tuple(int(v) for v in c.split(":")) for c in (sys.argv[3].split(",") if len(sys.argv) > 3 else ["40:128", "43:128", "46:128", "43:64", "43:256"])]
for nrc, ring in combos:
    os.environ["GPE_FUSED_NRC"] = str(nrc); os.environ["GPE_RING_TILES"] = str(ring)
    print(json.dumps({"nrc": nrc, "ring_tiles": ring, "fused_ms": timed()}), flush=True)
