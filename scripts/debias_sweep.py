"""Sweep the accumulator de-bias factor (GPE_TC_DEBIAS_FWD / _REV: the epilogues multiply what they read from tensor memory by
1 + beta 2^-24) and print the errors of the tensor path against the fp64 oracle: forward jets, the gradient of a random linear
functional of the jets, and the loss gradient of the GPE2D golden."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import numpy as np, torch
import gpe_b200
from gpe_b200 import ops
from oracle import jet_oracle as jo, pinn_oracle as po
from helpers import rel_err, load_weights, flat, model_grads
from test_jet_kernels import make_net, points, oracle_h, hbar_full

cases = []
for (d, W, L, jet, n) in [(2, 128, 4, 3, 3000), (2, 100, 3, 3, 1500), (1, 128, 4, 1, 2049)]:
    Ws, bs = make_net(d, W, L, seed=7 + W + L, act=0)
    x = points(n, d, seed=n)
    ur, gr, Hr = jo.jets(x, Ws, bs, 0)
    gen = torch.Generator().manual_seed(99)
    H = ops.jet_hess(d, jet)
    ub, gb, hb = torch.randn(n, generator=gen) * 0.3 + 1, torch.randn((n, d), generator=gen), torch.randn((n, H), generator=gen)
    rW, rb = jo.jet_functional_param_grads(x, Ws, bs, 0, ub, gb, hbar_full(hb, d, jet, ops))
    ref = np.concatenate([t.numpy().ravel() for t in rW + rb])
    cases.append((d, W, L, jet, n, [w.cuda() for w in Ws], [b.cuda() for b in bs], x.cuda(), ur, gr, oracle_h(Hr, d, jet, ops), ub.cuda(), gb.cuda(), hb.cuda(), ref))
g = np.load(os.path.join(ROOT, "tests", "golden", "g2d_g1_w128.npz"))
layers = [int(v) for v in g["layers"]]
refm = load_weights(po.GPE2D(layers, g=float(g["g"]), potential="harmonic"), g).double()
Xr = torch.tensor(g["X"], dtype=torch.float64, requires_grad=True)
total_r = po.gpe2d_pointwise_terms(refm, Xr, torch.tensor(g["Xb"], dtype=torch.float64))[0]
gref = flat(model_grads(refm, total_r))


def run(bf, br):
    os.environ["GPE_TC_DEBIAS_FWD"], os.environ["GPE_TC_DEBIAS_REV"] = str(bf), str(br)
    out = {"beta_fwd": bf, "beta_rev": br}
    for (d, W, L, jet, n, Ws, bs, x, ur, gr, hr, ub, gb, hb, ref) in cases:
        u, gg, h = ops.mlp_jet_fwd(x, Ws, bs, 0, jet)
        dWs, dbs = ops.mlp_jet_bwd(x, Ws, bs, 0, jet, ub, gb, hb)
        got = np.concatenate([t.cpu().numpy().ravel() for t in dWs + dbs])
        out[f"d{d}W{W}L{L}"] = [float(f"{v:.3g}") for v in (rel_err(u.cpu(), ur), rel_err(gg.cpu(), gr), rel_err(h.cpu(), hr),
                                                             float(np.linalg.norm(got - ref) / np.linalg.norm(ref)))]
    model = load_weights(gpe_b200.GPE2D(layers, g=float(g["g"]), potential="harmonic", semantics="pointwise"), g).cuda()
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    total = model.loss(X, torch.tensor(g["Xb"], dtype=torch.float32, device="cuda"), None)
    out["golden_grad"] = float(f"{rel_err(flat(model_grads(model, total)), gref):.3g}")
    print(json.dumps(out), flush=True)


betas = [float(v) for v in (sys.argv[1].split(",") if len(sys.argv) > 1 else "0,8,16,24,32,48".split(","))]
for b in betas:
    run(b, 0.0)
for b in betas[1:]:
    run(0.0, b)
for b in betas[1:]:
    run(b, b)
