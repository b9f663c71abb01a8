"""Large-n agreement of the tensor-path and FFMA-path gradients (accumulation error over many points)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import numpy as np, torch
from gpe_b200 import ops
from helpers import rel_err
from test_jet_kernels import make_net
for n in (4096, 65536, 1 << 20, 1 << 22):
    d, W, L, jet = 2, 128, 4, 3
    Ws, bs = make_net(d, W, L, seed=1, act=0)
    Ws = [w.cuda() for w in Ws]; bs = [b.cuda() for b in bs]
    torch.manual_seed(n)
    x = torch.rand(n, d, device="cuda") * 4 - 2
    for mode in ("signed", "positive"):
        ub = torch.randn(n, device="cuda"); gb = torch.randn(n, d, device="cuda"); hb = torch.randn(n, 1, device="cuda")
        if mode == "positive":
            ub, gb, hb = ub.abs(), gb.abs(), hb.abs()
        res = {}
        for name, flag in (("tensor", True), ("ffma", False)):
            ops.set_tensor_path(flag)
            dWs, dbs = ops.mlp_jet_bwd(x, Ws, bs, 0, jet, ub, gb, hb)
            res[name] = [t.double().cpu().numpy() for t in dWs + dbs]
        # fp64 reference of the same functional by torch autograd on chunks is too slow at 4M; the FFMA path
        # (fp32 FMA chains per CTA, fp64 cross-CTA reduction) is the comparison here
        per = " ".join(f"{rel_err(a, b):.1e}" for a, b in zip(res["tensor"], res["ffma"]))
        tot = rel_err(np.concatenate([a.ravel() for a in res["tensor"]]), np.concatenate([a.ravel() for a in res["ffma"]]))
        print(f"n={n} {mode}: tensor vs ffma total {tot:.2e} | per tensor {per}", flush=True)
ops.set_tensor_path(True)
