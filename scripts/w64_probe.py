import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import torch
from gpe_b200 import ops
from test_jet_kernels import make_net
for d, jet in ((1, 1), (2, 3)):
    for n in (4000, 65536, 1 << 20):
        for W in (64, 72):
            Ws, bs = make_net(d, W, 3, seed=1, act=0)
            Ws = [w.cuda() for w in Ws]; bs = [b.cuda() for b in bs]
            x = torch.rand(n, d, device="cuda") * 4 - 2
            H = ops.jet_hess(d, jet)
            ub = torch.randn(n, device="cuda"); gb = torch.randn(n, d, device="cuda"); hb = torch.randn(n, H, device="cuda")
            for _ in range(3):
                ops.mlp_jet_fwd(x, Ws, bs, 0, jet); ops.mlp_jet_bwd(x, Ws, bs, 0, jet, ub, gb, hb)
            a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            torch.cuda.synchronize(); a.record()
            for _ in range(10): ops.mlp_jet_fwd(x, Ws, bs, 0, jet)
            b.record()
            for _ in range(10): ops.mlp_jet_bwd(x, Ws, bs, 0, jet, ub, gb, hb)
            c.record(); torch.cuda.synchronize()
            print(f"d{d} n{n} W{W} ({'tensor' if W > 64 else 'ffma'}): fwd {a.elapsed_time(b) / 10:.3f} ms  bwd {b.elapsed_time(c) / 10:.3f} ms", flush=True)
