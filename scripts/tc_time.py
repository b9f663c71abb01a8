"""One forward + one backward call at the config-4 shape (for ncu launch lists / per-kernel timing)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import torch
from gpe_b200 import ops
from test_jet_kernels import make_net
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
d, W, L, jet = 2, 128, 4, 3
Ws, bs = make_net(d, W, L, seed=1, act=0)
Ws = [w.cuda() for w in Ws]; bs = [b.cuda() for b in bs]
x = torch.rand(n, d, device="cuda") * 4 - 2
ub = torch.randn(n, device="cuda"); gb = torch.randn(n, d, device="cuda"); hb = torch.randn(n, 1, device="cuda")
for _ in range(2):
    ops.mlp_jet_fwd(x, Ws, bs, 0, jet)
    ops.mlp_jet_bwd(x, Ws, bs, 0, jet, ub, gb, hb)
torch.cuda.synchronize()
print("ok")
