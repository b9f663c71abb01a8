"""Gradient / loss error of the tensor path against the fp64 oracle for the GPE2D golden (the case test_g2d_pointwise_vs_oracle
holds to 1e-5), printed instead of asserted -- for A/B builds of the library (csrc/build_variant.sh)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import numpy as np, torch
import gpe_b200
from oracle import pinn_oracle as po
from helpers import load_weights, flat, rel_err, model_grads

out = {}
for name in ("g2d_g1_w128",):
    g = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    layers = [int(v) for v in g["layers"]]
    ref = load_weights(po.GPE2D(layers, g=float(g["g"]), potential="harmonic"), g).double()
    Xr = torch.tensor(g["X"], dtype=torch.float64, requires_grad=True)
    total_r, bc_r, pde_r, riesz_r, mu_r = po.gpe2d_pointwise_terms(ref, Xr, torch.tensor(g["Xb"], dtype=torch.float64))
    gr = flat(model_grads(ref, total_r))
    model = load_weights(gpe_b200.GPE2D(layers, g=float(g["g"]), potential="harmonic", semantics="pointwise"), g).cuda()
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    xb = torch.tensor(g["Xb"], dtype=torch.float32, device="cuda")
    total = model.loss(X, xb, None)
    pde, res, mu = model.pde_loss(X, model.forward(X))
    gm = flat(model_grads(model, total))
    out[name] = {"grad": rel_err(gm, gr), "total": abs(float(total) - float(total_r)) / abs(float(total_r)),
                 "pde": abs(float(pde) - float(pde_r)) / abs(float(pde_r)), "mu": abs(float(mu) - float(mu_r)) / abs(float(mu_r))}
print(json.dumps({k: {a: float(f"{b:.3g}") for a, b in v.items()} for k, v in out.items()}))
