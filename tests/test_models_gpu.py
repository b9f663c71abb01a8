"""GPU parity of the drop-in modules (gpe_b200.models) against the golden vectors of the REAL reference
classes (tests/golden/*.npz, float64 runs) and against the oracle run on the same inputs.

The loop bodies are the oracle's own (`root_train_terms`, `final_train_terms`, ...): they only use the
reference method names, so the same function drives the oracle class on the CPU and the B200 class on
the GPU.  Tolerances (north-star): loss terms and parameter gradients 1e-5 relative in fp32 -- measured
here against the float64 goldens, so the bound includes the kernels' own fp32 rounding."""
import os

import numpy as np
import pytest
import torch

from oracle import pinn_oracle as po
from helpers import scalar, load_weights, golden_grads, flat, rel_err, model_grads, VARIANT_GOLDENS, variant_model_and_args, \
    OPT_VARIANT_GOLDENS, opt_variant_model_and_kwargs

pytestmark = pytest.mark.gpu
# north-star: loss terms AND parameter gradients within 1e-5 relative (measured: 1-4e-6).  GPE_TEST_TOL only ever TIGHTENS
# the bar (margin probes of kernel variants: which goldens sit close to it)
TOL = min(1e-5, float(os.environ.get("GPE_TEST_TOL", "1e-5")))
GTOL = TOL


def close(a, b, tol=TOL):
    a, b = scalar(a), scalar(b)
    return abs(a - b) <= tol * max(abs(b), 1e-30)


def bc_close(bc, ref):
    """boundary_loss = mean(u_b^2) at two points where u_b ~ 1e-3 is a cancelling sum of O(1) hidden activations: its
    fp32 value carries an ABSOLUTE error of ~eps(1) = 6e-8, i.e. 2 u_b 6e-8 ~ 5e-10 on the term, whatever evaluates it
    (the reference's own float32 run differs from its float64 run by as much)."""
    return abs(scalar(bc) - scalar(ref)) <= TOL * abs(scalar(ref)) + 5e-10


def norm_close(nrm, ref):
    """normalization_loss = (S - 1)^2 with S = sum(u^2) dx ~ 1 is the square of an fp32 cancellation: compare S itself
    (|S - 1| = sqrt of the term) to 1e-6 instead of its squared residual to a loose relative bound."""
    a, b = abs(scalar(nrm)) ** 0.5, abs(scalar(ref)) ** 0.5
    return abs(a - b) <= 1e-6 * (1.0 + b)


def cuda_model(cls, gold, *args, prefix="w.", **kw):
    return load_weights(cls(*args, **kw), gold, prefix=prefix).cuda()


@pytest.mark.parametrize("name", ["root_m0_p2", "root_m1_p3", "root_m2_p3"])
def test_root(golden, name):
    import gpe_b200
    g = golden(name)
    model = cuda_model(gpe_b200.RootGPE1D, g, [int(v) for v in g["layers"]], mode=int(g["mode"]), gamma=float(g["gamma"]),
                       power=int(g["p"]))
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    lb, ub = float(g["lb"]), float(g["ub"])
    bp = torch.tensor([[lb], [ub]], device="cuda")
    total, pde, bc, nrm, sym, mu = po.root_train_terms(model, X, bp, torch.zeros((2, 1), device="cuda"), float(g["dx"]),
                                                       float(g["gamma"]), int(g["p"]), lb, ub)
    for key, val in dict(pde=pde, bc=bc, norm=nrm, sym=sym, total=total, mu=mu).items():
        assert close(val, g[f"f64.{key}"]), (key, float(val), float(g[f"f64.{key}"]))
    _, res, _, full = model.pde_loss(int(g["p"]), X, model.forward(X), float(g["gamma"]))
    assert rel_err(full.detach().cpu().numpy().ravel(), g["f64.u"]) < TOL
    assert rel_err(res.detach().cpu().numpy().ravel(), g["f64.residual"]) < 5 * TOL
    assert rel_err(flat(model_grads(model, total)), flat(golden_grads(g, "f64.grads"))) < GTOL


@pytest.mark.parametrize("name", ["paper_m0_g10_p3", "paper_m1_g20_p2"])
def test_paper_notebook(golden, name):
    """gpe_b200.PaperGPE1D (NBP/Gross_Pitaevskii_1D_Harmonic.ipynb cell 6) with the loop body of cell 8:97-145."""
    import gpe_b200
    g = golden(name)
    model = cuda_model(gpe_b200.PaperGPE1D, g, [int(v) for v in g["layers"]], mode=int(g["mode"]), gamma=float(g["gamma"]))
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    lb, ub = float(g["lb"]), float(g["ub"])
    bp = torch.tensor([[lb], [ub]], device="cuda")
    gamma, p = float(g["gamma"]), int(g["p"])
    total, phys, bc, nrm, sym, lam = po.paper_train_terms(model, X, bp, torch.zeros((2, 1), device="cuda"), float(g["dx"]),
                                                          gamma, p, lb, ub)
    e, lam_e, _ = model.riesz_loss(X, model.forward(X), gamma, p)
    pde, res, mu, u = model.pde_loss(X, model.forward(X), gamma, p)
    for key, val in dict(pde=pde, riesz=e, lam_riesz=lam_e, mu=mu, bc=bc, sym=sym, total=total).items():
        assert close(val, g[f"f64.{key}"], TOL), (key, float(val), float(g[f"f64.{key}"]))
    assert close(lam, g["f64.lam_riesz"] if int(g["mode"]) == 0 else g["f64.mu"], TOL)
    assert norm_close(nrm, g["f64.norm"]), (float(nrm), float(g["f64.norm"]))
    assert rel_err(u.detach().cpu().numpy()[:, 0], g["f64.u"]) < TOL
    assert rel_err(res.detach().cpu().numpy()[:, 0], g["f64.residual"]) < 5 * TOL
    assert rel_err(flat(model_grads(model, total)), flat(golden_grads(g, "f64.grads"))) < GTOL


@pytest.mark.parametrize("name,kind", [("final_harmonic_g20_p3_m1", "harmonic"), ("final_harmonic_g5_p4_m2", "harmonic"),
                                       ("final_box_g10_p3_m0", "box"), ("final_gravity_g10_p3_m1", "gravity_well"),
                                       ("final_box_vanilla_g10_p3_m0", "box"), ("final_box_option_g5_p3_m1", "box"),
                                       ("final_neg_gamma_shipped_m1_gm8", "harmonic"),
                                       ("final_curriculum_shipped_m0_g20", "harmonic"), ("final_curriculum_shipped_m1_g60", "harmonic")])
def test_final_random(golden, name, kind):
    import gpe_b200
    g = golden(name)
    opts = {k: bool(g[k]) for k in ("use_perturbation", "enforce_boundary") if k in g}   # plot_box_potential_at_ground_state.py
    model = cuda_model(gpe_b200.FinalGPE1D, g, [int(v) for v in g["layers"]], mode=int(g["mode"]), gamma=float(g["gamma"]),
                       potential=kind, **opts)
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    lb, ub = float(g["lb"]), float(g["ub"])
    bp = torch.tensor([[lb], [ub]], device="cuda")
    total, pde, bc, nrm, mu = po.final_train_terms(model, X, bp, torch.zeros((2, 1), device="cuda"), float(g["dx"]),
                                                   float(g["gamma"]), int(g["p"]), float(g["perturb_const"]), float(g["const"]))
    for key, val in dict(pde=pde, total=total, mu=mu).items():
        assert close(val, g[f"f64.{key}"], TOL), (key, float(val), float(g[f"f64.{key}"]))
    # (sum u^2 dx - 1)^2 of a nearly normalised state: squares an O(1e-4) fp32 difference
    assert norm_close(nrm, g["f64.norm"]), (float(nrm), float(g["f64.norm"]))
    assert bc_close(bc, g["f64.bc"]), (float(bc), float(g["f64.bc"]))
    assert rel_err(flat(model_grads(model, total)), flat(golden_grads(g, "f64.grads"))) < GTOL


B2G_GOLDENS = ["box_to_gaussian_res_g10_p3_m0", "box_to_gaussian_res_g4_p4_m2", "box_to_gaussian_plain_g5_p3_m1"]


def _b2g_check(model, g, tol, gtol):
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    lb, ub = float(g["lb"]), float(g["ub"])
    bp = torch.tensor([[lb], [ub]], device="cuda")
    total, pde, bc, nrm, mu = po.final_train_terms(model, X, bp, torch.zeros((2, 1), device="cuda"), float(g["dx"]),
                                                   float(g["gamma"]), int(g["p"]), float(g["perturb_const"]), float(g["const"]))
    for key, val in dict(pde=pde, total=total, mu=mu).items():
        assert close(val, g[f"f64.{key}"], tol), (key, float(val), float(g[f"f64.{key}"]))
    assert norm_close(nrm, g["f64.norm"]), (float(nrm), float(g["f64.norm"]))
    assert abs(float(bc) - float(g["f64.bc"])) <= tol * abs(float(g["f64.bc"])) + 1e-12
    assert rel_err(flat(model_grads(model, total)), flat(golden_grads(g, "f64.grads"))) < gtol


@pytest.mark.parametrize("name", B2G_GOLDENS)
def test_box_to_gaussian_residual_blocks(golden, name):
    """the ResidualBlock network of R/box_to_gaussian_pinn_simulation.py through the fused path
    (GPE_ACT_RESBLOCK_TANH kernels + moment epilogue), against the float64 run of the real class."""
    import gpe_b200
    g = golden(name)
    model = cuda_model(gpe_b200.BoxToGaussianGPE1D, g, [int(v) for v in g["layers"]], mode=int(g["mode"]),
                       gamma=float(g["gamma"]), use_residual="_res_" in name)
    assert ("network.2.lin1.weight" in model.state_dict()) == ("_res_" in name)
    _b2g_check(model, g, TOL, GTOL)


def test_compat_box_to_gaussian_residual_blocks(golden):
    """... and through the compat path: the oracle's (= reference's) class with only forward() rerouted; its own
    nested autograd.grad calls, which also differentiate the torch box-sine base, run on the kernel's jets."""
    from gpe_b200.models import patch
    g = golden("box_to_gaussian_res_g10_p3_m0")
    model = patch(load_weights(po.BoxToGaussianGPE1D([int(v) for v in g["layers"]], mode=int(g["mode"]),
                                                     gamma=float(g["gamma"])), g).cuda())
    _b2g_check(model, g, TOL, GTOL)


@pytest.mark.parametrize("kind", ["harmonic", "box", "gravity_well"])
def test_final_fixture_mu_table(golden, kind):
    """Shipped trained weights: the fused path must reproduce the pickled mu_table;
    converged mu within 1e-4 (north-star) -- in fact ~1e-6."""
    import gpe_b200
    g = golden(f"final_{kind.split('_')[0]}_fixture")
    lb, ub = float(g["lb"]), float(g["ub"])
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    bp = torch.tensor([[lb], [ub]], device="cuda")
    for mode in range(6):
        model = cuda_model(gpe_b200.FinalGPE1D, g, [int(v) for v in g[f"m{mode}.layers"]], prefix=f"m{mode}.w.", mode=mode,
                           gamma=0.0, potential=kind)
        total, pde, bc, nrm, mu = po.final_train_terms(model, X, bp, torch.zeros((2, 1), device="cuda"), float(g["dx"]),
                                                       0.0, 3, 0.01, float(g[f"m{mode}.const"]))
        assert abs(float(mu) - float(g[f"m{mode}.mu_table"])) < TOL * abs(float(g[f"m{mode}.mu_table"])), mode
        assert abs(float(mu) - float(g[f"m{mode}.f64.mu"])) < TOL * abs(float(g[f"m{mode}.f64.mu"])), mode
        # converged losses are ~1e-6 differences of O(1) numbers: fp32 agreement is loose by nature
        for key, val in dict(pde=pde, total=total).items():
            ref = float(g[f"m{mode}.f64.{key}"])
            assert abs(float(val) - ref) <= 0.3 * abs(ref) + 1e-9, (mode, key, float(val), ref)


@pytest.mark.parametrize("name", VARIANT_GOLDENS)
def test_beta_sweep_variants(golden, name):
    """gpe_b200.VaryPotentialGPE1D / PerturbGammaBetaGPE1D (R/vary_potential_parameter_*.py,
    R/test_perturbing_gamma_and_beta.py) against the float64 runs of the real classes."""
    import gpe_b200
    g = golden(name)
    model, pde_args = variant_model_and_args(gpe_b200, g)
    model = model.cuda()
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    lb, ub = float(g["lb"]), float(g["ub"])
    bp = torch.tensor([[lb], [ub]], device="cuda")
    total, pde, bc, nrm, mu = po.variant_train_terms(model, X, bp, torch.zeros((2, 1), device="cuda"), float(g["dx"]),
                                                     pde_args, float(g["perturb_const"]), float(g["const"]))
    shipped = "shipped" in name
    assert close(mu, g["f64.mu"], TOL), (float(mu), float(g["f64.mu"]))
    if shipped:      # converged weights: losses are ~1e-6 differences of O(1) numbers (see test_final_fixture_mu_table)
        for key, val in dict(pde=pde, total=total).items():
            ref = float(g[f"f64.{key}"])
            assert abs(float(val) - ref) <= 0.3 * abs(ref) + 1e-9, (key, float(val), ref)
        assert abs(float(mu) - float(g["mu_logged"])) < 2e-3 * abs(float(g["mu_logged"]))
        return
    for key, val in dict(pde=pde, total=total).items():
        assert close(val, g[f"f64.{key}"], TOL), (key, float(val), float(g[f"f64.{key}"]))
    assert norm_close(nrm, g["f64.norm"]), (float(nrm), float(g["f64.norm"]))
    assert bc_close(bc, g["f64.bc"]), (float(bc), float(g["f64.bc"]))
    assert rel_err(flat(model_grads(model, total)), flat(golden_grads(g, "f64.grads"))) < GTOL


@pytest.mark.parametrize("name", OPT_VARIANT_GOLDENS)
def test_opt_variants(golden, name):
    """gpe_b200.SelfAdaptiveGPE1D / PowersGPE1D / ModesGPE1D against the float64 runs of the real classes
    (_Self_Adaptive.py, _and_Nonlinearity_Powers.py, _Different_Modes.py)."""
    import gpe_b200
    g = golden(name)
    model, extra = opt_variant_model_and_kwargs(gpe_b200, g)
    model = model.cuda()
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    lb, ub = torch.tensor(float(g["lb"]), device="cuda"), torch.tensor(float(g["ub"]), device="cuda")
    bp = torch.stack([lb, ub]).view(2, 1)
    eta = float(g["eta"])
    ret = model.total_loss(X, bp, torch.zeros((2, 1), device="cuda"), eta, lb, ub, list(g["weights"]), "harmonic", **extra)
    _, _, mu = model.pde_loss(X, model.forward(X), eta, "harmonic", **extra)
    got = dict(total=ret[0], data=ret[1], riesz=ret[2], pde=ret[3], mu=mu)
    if str(g["kind"]) == "self_adaptive":
        got["sym"] = ret[5]
        assert rel_err(ret[6].detach().cpu().numpy(), g["f64.alpha"]) < TOL
    for key, val in got.items():
        assert close(val, g[f"f64.{key}"], TOL), (key, float(val), float(g[f"f64.{key}"]))
    assert abs(float(ret[4]) - float(g["f64.norm"])) <= 1e-4 * abs(float(g["f64.norm"])) + 1e-9
    assert rel_err(flat(model_grads(model, ret[0])), flat(golden_grads(g, "f64.grads"))) < GTOL


@pytest.mark.parametrize("name", ["opt_m0_eta10", "opt_m1_eta50"])
def test_opt(golden, name):
    import gpe_b200
    g = golden(name)
    model = cuda_model(gpe_b200.OptGPE1D, g, [int(v) for v in g["layers"]], mode=int(g["mode"]))
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    lb, ub = torch.tensor(float(g["lb"])), torch.tensor(float(g["ub"]))
    bp = torch.stack([lb, ub]).reshape(2, 1).cuda()
    w = [float(v) for v in g["weights"]]
    total, data, riesz, pde, norm = model.total_loss(X, bp, torch.zeros((2, 1), device="cuda"), float(g["eta"]), lb, ub, w, "harmonic")
    five = model.last_weighted_sum
    sym = model.symmetry_loss(X, lb, ub)
    _, res, mu = model.pde_loss(X, model.forward(X), float(g["eta"]), "harmonic")
    assert tuple(res.shape) == (X.shape[0], 1)
    for key, val in dict(total=total, data=data, riesz=riesz, pde=pde, norm=norm, sym=sym, mu=mu, five=five).items():
        assert close(val, g[f"f64.{key}"]), (key, float(val), float(g[f"f64.{key}"]))
    assert rel_err(flat(model_grads(model, five)), flat(golden_grads(g, "f64.grads_five"))) < GTOL
    total2 = model.total_loss(X, bp, torch.zeros((2, 1), device="cuda"), float(g["eta"]), lb, ub, w, "harmonic")[0]
    assert rel_err(flat(model_grads(model, total2)), flat(golden_grads(g, "f64.grads"))) < GTOL


@pytest.mark.parametrize("name", ["g2d_g100", "g2d_g1_w128"])
def test_g2d_reference_semantics(golden, name):
    import gpe_b200
    g = golden(name)
    model = cuda_model(gpe_b200.GPE2D, g, [int(v) for v in g["layers"]], g=float(g["g"]))
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    xb = torch.tensor(g["Xb"], dtype=torch.float32, device="cuda")
    total = model.loss(X, xb, None)
    pde, res, mu = model.pde_loss(X, model.forward(X))
    assert tuple(res.shape) == tuple(int(v) for v in g["f64.res_shape"])       # the [N,N] broadcast
    for key, val in dict(total=total, pde=pde, mu=mu, bc=model.boundary_loss(xb, None),
                         riesz=model.riesz_loss(model.forward(X), X)).items():
        assert close(val, g[f"f64.{key}"]), (key, float(val), float(g[f"f64.{key}"]))
    assert rel_err(flat(model_grads(model, total)), flat(golden_grads(g, "f64.grads"))) < GTOL


def test_g2d_pointwise_vs_oracle(golden):
    import gpe_b200
    g = golden("g2d_g1_w128")
    layers = [int(v) for v in g["layers"]]
    ref = load_weights(po.GPE2D(layers, g=float(g["g"]), potential="harmonic"), g).double()
    model = cuda_model(gpe_b200.GPE2D, g, layers, g=float(g["g"]), potential="harmonic", semantics="pointwise")
    Xr = torch.tensor(g["X"], dtype=torch.float64, requires_grad=True)
    total_r, bc_r, pde_r, riesz_r, mu_r = po.gpe2d_pointwise_terms(ref, Xr, torch.tensor(g["Xb"], dtype=torch.float64))
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    xb = torch.tensor(g["Xb"], dtype=torch.float32, device="cuda")
    total = model.loss(X, xb, None)
    pde, res, mu = model.pde_loss(X, model.forward(X))
    assert tuple(res.shape) == (X.shape[0], 1)
    for a, b in ((total, total_r), (pde, pde_r), (mu, mu_r)):
        assert close(a, b), (float(a), float(b))
    assert rel_err(flat(model_grads(model, total)), flat(model_grads(ref, total_r))) < GTOL


@pytest.mark.parametrize("name", ["helmholtz_leaky", "helmholtz_tanh"])
def test_helmholtz(golden, name):
    import gpe_b200
    g = golden(name)
    model = cuda_model(gpe_b200.Helmholtz2D, g, [int(v) for v in g["layers"]], activation=str(g["act"]))
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda")
    xb, ub = torch.tensor(g["Xb"], dtype=torch.float32, device="cuda"), torch.tensor(g["ub"], dtype=torch.float32, device="cuda")
    k = float(g["k"])
    total = model.compute_loss(X, xb, ub, k)
    for key, val in dict(pde=model.loss_pde(X, k), data=model.loss_data(xb, ub), total=total).items():
        assert close(val, g[f"f64.{key}"]), (key, float(val), float(g[f"f64.{key}"]))
    assert rel_err(flat(model_grads(model, total)), flat(golden_grads(g, "f64.grads"))) < GTOL
    # pointwise semantics against the oracle's intended-maths method
    ref = load_weights(po.Helmholtz2D([int(v) for v in g["layers"]], activation=str(g["act"])), g).double()
    model.semantics = "pointwise"
    assert close(model.loss_pde(X, k), ref.loss_pde_pointwise(torch.tensor(g["X"], dtype=torch.float64), k))


# ------------------------------------------------------------------------------------------------
# compat path: the ORACLE's (= reference's) own loss code, nested autograd.grad included, on top of
# the jet node -- only forward() is rerouted (gpe_b200.models.patch)
# ------------------------------------------------------------------------------------------------
def test_compat_root(golden):
    from gpe_b200.models import patch
    g = golden("root_m1_p3")
    model = patch(load_weights(po.RootGPE1D([int(v) for v in g["layers"]], mode=int(g["mode"]), gamma=float(g["gamma"]),
                                            power=int(g["p"])), g).cuda())
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    lb, ub = float(g["lb"]), float(g["ub"])
    bp = torch.tensor([[lb], [ub]], device="cuda")
    total, pde, bc, nrm, sym, mu = po.root_train_terms(model, X, bp, torch.zeros((2, 1), device="cuda"), float(g["dx"]),
                                                       float(g["gamma"]), int(g["p"]), lb, ub)
    for key, val in dict(pde=pde, bc=bc, norm=nrm, sym=sym, total=total, mu=mu).items():
        assert close(val, g[f"f64.{key}"], TOL), (key, float(val), float(g[f"f64.{key}"]))
    assert rel_err(flat(model_grads(model, total)), flat(golden_grads(g, "f64.grads"))) < GTOL


def test_compat_g2d_and_helmholtz(golden):
    from gpe_b200.models import patch
    g = golden("g2d_g100")
    model = patch(load_weights(po.GPE2D([int(v) for v in g["layers"]], g=float(g["g"])), g).cuda())
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    xb = torch.tensor(g["Xb"], dtype=torch.float32, device="cuda")
    total = model.loss(X, xb, None)
    assert close(total, g["f64.total"], TOL)
    assert rel_err(flat(model_grads(model, total)), flat(golden_grads(g, "f64.grads"))) < GTOL
    g = golden("helmholtz_tanh")
    model = patch(load_weights(po.Helmholtz2D([int(v) for v in g["layers"]], activation="tanh"), g).cuda())
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda")
    xb, ub = torch.tensor(g["Xb"], dtype=torch.float32, device="cuda"), torch.tensor(g["ub"], dtype=torch.float32, device="cuda")
    total = model.compute_loss(X, xb, ub, float(g["k"]))
    assert close(total, g["f64.total"], TOL)
    assert rel_err(flat(model_grads(model, total)), flat(golden_grads(g, "f64.grads"))) < GTOL


def _relobralo_check(model, g, tol_scale=1.0):
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    xb = torch.tensor(g["Xb"], dtype=torch.float32, device="cuda")
    bv = torch.zeros((xb.shape[0], 1), device="cuda")
    eta = float(g["eta"])
    torch.manual_seed(77)
    for c in range(int(g["calls"])):
        total, data, riesz, pde, norm = model.total_loss(X, xb, bv, eta)
        assert close(total, g[f"f64.c{c}.total"], TOL * tol_scale), (c, float(total), float(g[f"f64.c{c}.total"]))
        assert rel_err(np.array(model.lambdas), g[f"f64.c{c}.lambdas"]) < TOL * tol_scale, c
    grads = model_grads(model, total)
    _, res, lam = model.pde_loss(X, model.forward(X), eta)
    for key, val in dict(data=data, riesz=riesz, pde=pde, sym=model.symmetry_loss(X), mu=lam).items():
        assert close(val, g[f"f64.{key}"], TOL * tol_scale), (key, float(val), float(g[f"f64.{key}"]))
    assert abs(float(norm) - float(g["f64.norm"])) <= 1e-4 * abs(float(g["f64.norm"])) + 1e-9
    assert rel_err(res.detach().cpu().numpy(), g["f64.residual"]) < 5 * TOL * tol_scale
    assert rel_err(flat(grads), flat(golden_grads(g, "f64.grads"))) < GTOL * tol_scale


def test_relobralo_2d(golden):
    """gpe_b200.ReLoBRaLoGPE2D (S/gross_pitaevskii_2D_ReLoBRaLo.py): FULL jets + one moment pass per residual column,
    three calls of the ReLoBRaLo-weighted total, against the float64 run of the real class."""
    import gpe_b200
    g = golden("g2d_relobralo_eta10")
    _relobralo_check(cuda_model(gpe_b200.ReLoBRaLoGPE2D, g, [int(v) for v in g["layers"]]), g)


def test_compat_relobralo_2d(golden):
    """The same file through patch(): its own loss code (Hessian.(1,1) by nested autograd.grad) on the jet node."""
    from gpe_b200.models import patch
    g = golden("g2d_relobralo_eta10")
    _relobralo_check(patch(load_weights(po.ReLoBRaLoGPE2D([int(v) for v in g["layers"]]), g).cuda()), g, tol_scale=1.0)


def test_g2d_minimal(golden):
    """S/gross_pitaevskii_2D_minimal.py:201-222 total_loss on gpe_b200.GPE2D."""
    import gpe_b200
    g = golden("g2d_minimal_g50")
    model = cuda_model(gpe_b200.GPE2D, g, [int(v) for v in g["layers"]], g=float(g["g"]))
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    xb = torch.tensor(g["Xb"], dtype=torch.float32, device="cuda")
    total = model.total_loss(X, xb, None)
    assert close(total, g["f64.total"], TOL)
    assert rel_err(flat(model_grads(model, total)), flat(golden_grads(g, "f64.grads"))) < GTOL


def test_compat_helmholtz_circle(golden):
    """S/helmholtz_2D_circle.py through patch(): the forward's min-max scaling (:89-94) is folded into the first
    layer, the reference's polar loss (:112-172; three nested autograd.grad calls, [N,N] tables) runs unchanged on
    the kernel's u, grad u and Hessian.  Truth: the oracle in float64 (pinned to the real class in
    test_oracle_golden.test_helmholtz_circle)."""
    from gpe_b200.models import patch
    g = golden("helmholtz_circle")
    layers = [int(v) for v in g["layers"]]
    ref = load_weights(po.HelmholtzCircle(layers, lb=g["lb"], ub=g["ub"]), g).double()
    X64 = torch.tensor(g["X"], dtype=torch.float64)
    want = ref.loss_PDE(X64)
    want_g = flat(model_grads(ref, want))
    model = patch(load_weights(po.HelmholtzCircle(layers, lb=g["lb"], ub=g["ub"]), g).cuda(), input_lb=g["lb"], input_ub=g["ub"])
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda")
    xg = X.clone().requires_grad_(True)
    u = model(xg)
    du = po.d_dx(u, xg)
    d2 = po.d_dx(du, xg)
    assert rel_err(u.detach().cpu().numpy(), g["f32.u"]) < 1e-5
    assert rel_err(du.detach().cpu().numpy(), g["f32.du"]) < 1e-5
    assert rel_err(d2.detach().cpu().numpy(), g["f32.d2"]) < 1e-5
    got = model.loss_PDE(X)
    assert close(got, want, TOL), (float(got), float(want))
    assert rel_err(flat(model_grads(model, got)), want_g) < GTOL


def test_compat_helmholtz_circle_2_and_3(golden):
    """S/helmholtz_2D_circle_2.py and _circle_3.py through patch(): the files' own polar losses run unchanged on the
    kernel's u, grad u and Hessian (FULL jet); truth = the oracle in float64, pinned to the real classes in
    test_oracle_golden.test_helmholtz_circle_2_and_3."""
    from gpe_b200.models import patch
    g = golden("helmholtz_circle_2")
    layers = [int(v) for v in g["layers"]]
    ref = load_weights(po.HelmholtzCircle2(layers, lb=g["lb"], ub=g["ub"]), g).double()
    want = ref.loss_PDE(torch.tensor(g["X"], dtype=torch.float64), 1, 1)
    want_g = flat(model_grads(ref, want))
    model = patch(load_weights(po.HelmholtzCircle2(layers, lb=g["lb"], ub=g["ub"]), g).cuda(), input_lb=g["lb"], input_ub=g["ub"])
    got = model.loss_PDE(torch.tensor(g["X"], dtype=torch.float32, device="cuda"), 1, 1)
    assert close(got, want, TOL), (float(got), float(want))
    assert rel_err(flat(model_grads(model, got)), want_g) < GTOL
    g = golden("helmholtz_circle_3")
    layers = [int(v) for v in g["layers"]]
    model = patch(load_weights(po.HelmholtzCircle3(layers), g).cuda())
    got = model.loss_pde(torch.tensor(g["X"], dtype=torch.float32, device="cuda"), int(g["n"]), float(g["k"]))
    assert close(got, g["f64.pde"], TOL), (float(got), float(g["f64.pde"]))
    assert rel_err(flat(model_grads(model, got)), flat(golden_grads(g, "f64.grads"))) < GTOL


def test_helmholtz_learnable_k(golden):
    """gpe_b200.LearnableKHelmholtz2D (notebook #18): loss terms, and the gradient w.r.t. adaptive_bc_scale, k and the
    network, against the float64 run of the real class."""
    import gpe_b200
    g = golden("helmholtz_learnable_k")
    model = cuda_model(gpe_b200.LearnableKHelmholtz2D, g, [int(v) for v in g["layers"]], list(g["ub"]), list(g["lb"]))
    assert [n for n, _ in model.named_parameters()][:3] == ["adaptive_bc_scale", "k", "layers.0.weight"]
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    xb = torch.tensor(g["Xb"], dtype=torch.float32, device="cuda")
    yb = torch.tensor(g["yb"], dtype=torch.float32, device="cuda")
    pred = model.forward(X)
    terms = dict(pde=model.pde_loss(X, pred), riesz=model.riesz_loss(pred, X), bc=model.loss_BC(xb, yb), total=model.loss(xb, yb, X))
    for key, val in terms.items():
        assert close(val, g[f"f64.{key}"], TOL), (key, float(val), float(g[f"f64.{key}"]))
    grads, ref = model_grads(model, terms["total"]), golden_grads(g, "f64.grads")
    assert rel_err(flat(grads), flat(ref)) < GTOL
    for i in (0, 1):                                   # the two scalar parameters on their own
        assert abs(float(grads[i]) - float(ref[i])) <= GTOL * abs(float(ref[i])), (i, float(grads[i]), float(ref[i]))


def test_compat_helmholtz_learnable_k(golden):
    """... and the oracle's (= notebook's) class with only forward() rerouted (input scaling folded by patch)."""
    from gpe_b200.models import patch
    g = golden("helmholtz_learnable_k")
    model = load_weights(po.LearnableKHelmholtz2D([int(v) for v in g["layers"]], list(g["ub"]), list(g["lb"])), g).cuda()
    model = patch(model, input_lb=list(g["lb"]), input_ub=list(g["ub"]))
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    xb = torch.tensor(g["Xb"], dtype=torch.float32, device="cuda")
    yb = torch.tensor(g["yb"], dtype=torch.float32, device="cuda")
    total = model.loss(xb, yb, X)
    assert close(total, g["f64.total"], TOL)
    assert rel_err(flat(model_grads(model, total)), flat(golden_grads(g, "f64.grads"))) < GTOL


def test_helmholtz2d_input_scaling():
    """gpe_b200.Helmholtz2D(lb=, ub=): fused loss_pde with the scaling folded into (W_0, b_0); gradients reach the
    unscaled leaves.  Against the fp64 oracle with the scaling applied as the reference writes it."""
    import gpe_b200
    torch.manual_seed(3)
    layers, lb, ub = [2, 128, 128, 128, 1], [-1.0, -2.0], [3.0, 1.5]
    ref = po.Helmholtz2D(layers, activation="tanh", lb=lb, ub=ub).double()
    gpu = gpe_b200.Helmholtz2D(layers, activation="tanh", lb=lb, ub=ub)
    gpu.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    gpu = gpu.cuda()
    X64 = torch.rand(500, 2, dtype=torch.float64) * torch.tensor([4.0, 3.5]) + torch.tensor(lb)
    k = float(np.sqrt(2.0) * np.pi)
    want = ref.loss_pde(X64, k)
    got = gpu.loss_pde(X64.float().cuda().requires_grad_(True), k)
    assert close(got, want, TOL), (float(got), float(want))
    assert rel_err(flat(model_grads(gpu, got)), flat(model_grads(ref, want))) < GTOL


def test_reverse_kernel_runs_once_per_step(golden):
    """The C++ node must not run the parameter-reverse kernel during the nested autograd.grad(u, x)
    visits (SURVEY 8b-A note ii): count launches through the profiler-free route of a call counter."""
    import gpe_b200
    from gpe_b200.models import patch
    g = golden("root_m0_p2")
    model = patch(load_weights(po.RootGPE1D([int(v) for v in g["layers"]], mode=0, gamma=1.0, power=2), g).cuda())
    X = torch.tensor(g["X"], dtype=torch.float32, device="cuda", requires_grad=True)
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA, torch.profiler.ProfilerActivity.CPU]) as prof:
        pde, _, _, _ = model.pde_loss(2, X, model.forward(X), 1.0)
        pde.backward()
        torch.cuda.synchronize()
    # FFMA kernels (jet_fwd_kernel / jet_bwd_kernel) or the tensor path's (jet_fwd*_tc_kernel with the stash / jet_rev*_tc_kernel)
    n_bwd = sum(1 for e in prof.events() if "jet_bwd_kernel" in e.name or ("jet_rev" in e.name and "tc_kernel" in e.name))
    n_fwd = sum(1 for e in prof.events() if "jet_fwd_kernel" in e.name or ("jet_fwd" in e.name and "tc_kernel" in e.name))
    assert n_fwd == 1 and n_bwd == 1, (n_fwd, n_bwd)


def test_cpu_tensor_raises():
    import gpe_b200
    model = gpe_b200.RootGPE1D([1, 16, 16, 1])
    with pytest.raises(RuntimeError):
        model.forward(torch.zeros(4, 1))


def test_helmholtz_reference_default_width_200():
    """The reference's own main (S/helmholtz_2D.py:378) uses [2,200,200,200,1] with LeakyReLU: hidden width padded to
    256 (FFMA kernels).  Loss (reference [N,N] semantics) and gradients against the fp64 oracle on the same weights."""
    import gpe_b200
    torch.manual_seed(0)
    layers = [2, 200, 200, 200, 1]
    ref = po.Helmholtz2D(layers).double()
    gpu = gpe_b200.Helmholtz2D(layers, semantics="reference")
    gpu.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    gpu = gpu.cuda()
    X = torch.rand(500, 2, dtype=torch.float64) * np.pi
    Xu, U = torch.rand(80, 2, dtype=torch.float64) * np.pi, torch.rand(80, 1, dtype=torch.float64)
    k = float(np.sqrt(2) * np.pi)
    lr = ref.compute_loss(X.clone().requires_grad_(True), Xu, U, k)
    lg = gpu.compute_loss(X.float().cuda().requires_grad_(True), Xu.float().cuda(), U.float().cuda(), k)
    assert close(lg, lr)
    gr = [p.grad.numpy() for p in ref.parameters()] if lr.backward() is None else None
    assert rel_err(flat(model_grads(gpu, lg)), flat(gr)) < GTOL
