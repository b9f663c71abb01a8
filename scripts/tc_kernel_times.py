"""Per-kernel times of the tensor-path backward at the config-4 shape (CUDA events on the launch stream via
gpe_tc_profile).  usage: tc_kernel_times.py [n_points] [reps]"""
import os, sys, json
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
for _ in range(2):
    ops.mlp_jet_fwd(x, Ws, bs, 0, jet)
    ops.mlp_jet_bwd(x, Ws, bs, 0, jet, ub, gb, hb)
torch.cuda.synchronize()
a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
a.record()
for _ in range(reps): ops.mlp_jet_fwd(x, Ws, bs, 0, jet)
b.record()
for _ in range(reps): ops.mlp_jet_bwd(x, Ws, bs, 0, jet, ub, gb, hb)
c.record(); torch.cuda.synchronize()
ops.tc_profile(True)
ops.mlp_jet_bwd(x, Ws, bs, 0, jet, ub, gb, hb)
t = ops.tc_timings(); ops.tc_profile(False)
print(json.dumps({"n": n, "env": {k: v for k, v in os.environ.items() if k.startswith("GPE_")}, "fwd_ms": a.elapsed_time(b) / reps,
                  "bwd_ms": b.elapsed_time(c) / reps, "ns_per_point": (a.elapsed_time(c)) / reps / n * 1e6, "kernels_ms": t}), flush=True)
