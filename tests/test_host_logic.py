"""CPU tests of the host side: the C-ABI library loads and exports every symbol include/gpe_b200.h
declares, fails loudly without a device, the moment algebra of gpe_b200.losses reproduces the oracle's
loss terms (with the moments restated in torch from the oracle's autograd jets -- test infrastructure),
the grad_fn scale tracer, and the world_size-2 data-parallel path over gloo."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from oracle import jet_oracle as jo
from oracle import pinn_oracle as po
from helpers import load_weights

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "gpe_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gpe_[a-z0-9_]+)\s*\(", src)))


def test_capi_loads_and_exports_every_declared_symbol():
    from gpe_b200 import _lib
    syms = declared_symbols()
    assert {"gpe_mlp_jet_fwd", "gpe_mlp_jet_bwd", "gpe_gram_moments", "gpe_gram_adjoints", "gpe_workspace_bytes",
            "gpe_last_error", "gpe_version", "gpe_ffma_probe"} <= set(syms)
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for s in syms:
        assert hasattr(lib, s), s
    assert set(_lib.EXPORTS) == set(syms)
    assert "sm_100a" in _lib.version()


def test_torch_ops_registered():
    import gpe_b200  # noqa: F401
    for name in ("mlp_jet_fwd", "mlp_jet_bwd", "mlp_jets", "gram_moments", "gram_adjoints", "ffma_probe"):
        assert hasattr(torch.ops.gpe, name)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-device failure mode")
def test_fails_loudly_without_a_device():
    from gpe_b200 import _lib
    import gpe_b200
    desc = _lib.MlpDesc(1, 3, 64, 0, 1, 2)
    fwd, bwd = ctypes.c_size_t(0), ctypes.c_size_t(0)
    rc = _lib.lib.gpe_workspace_bytes(ctypes.byref(desc), 100, ctypes.byref(fwd), ctypes.byref(bwd))
    assert rc < 0 and b"CPU fallback" in _lib.lib.gpe_last_error()
    model = gpe_b200.RootGPE1D([1, 16, 16, 1])
    with pytest.raises(RuntimeError, match="no CPU path"):
        model.forward(torch.zeros(4, 1))
    with pytest.raises(RuntimeError, match="CUDA tensor|CPU. backend"):
        gpe_b200.ops.mlp_jet_fwd(torch.zeros(4, 1), [torch.zeros(8, 1), torch.zeros(1, 8)], [torch.zeros(8), torch.zeros(1)], 0, 1)


# ---------------------------------------------------------------------------------------------
# moment algebra
# ---------------------------------------------------------------------------------------------
def torch_moments(u, gx, gy, lap, V, nl, q, n0=None, nm=None, sign=1):
    """Restatement of gram_moments_kernel in float64 torch (layout of include/gpe_b200.h)."""
    from gpe_b200 import _lib
    f = [torch.ones_like(u), u, lap, V, V * u, nl, q]
    m = torch.zeros(_lib.NMOM, dtype=torch.float64)
    for i in range(7):
        for j in range(i, 7):
            m[_lib.mom(i, j)] = (f[i] * f[j]).sum()
    m[_lib.MOM_GRAD2] = (gx * gx + gy * gy).sum()
    if n0 is not None:
        m[_lib.MOM_NN2] = (n0 * n0).sum()
    if nm is not None:
        m[_lib.MOM_SYM2] = ((n0 - sign * nm) ** 2).sum()
    return m


def oracle_jets(model, X):
    lins = po.linear_layers(model)
    act = {"Tanh": 0, "ShiftedTanh": 1, "LeakyReLU": 2}[[type(m).__name__ for m in model.modules()
                                                         if type(m).__name__ in ("Tanh", "ShiftedTanh", "LeakyReLU")][0]]
    return jo.jets(X, [l.weight for l in lins], [l.bias for l in lins], act)


@pytest.mark.parametrize("name", ["g2d_g100", "g2d_g1_w128"])
def test_g2d_moment_algebra(golden, name):
    from gpe_b200 import losses as L
    from gpe_b200._lib import F_U
    g = golden(name)
    model = load_weights(po.GPE2D([int(v) for v in g["layers"]], g=float(g["g"])), g).double()
    X = torch.tensor(g["X"], dtype=torch.float64)
    u, gr, H = oracle_jets(model, X)
    V = model.compute_potential(X)
    M = L.Moments(torch_moments(u, gr[:, 0], gr[:, 1], H[:, 0, 0] + H[:, 1, 1], V, u ** 3, torch.zeros_like(u)))
    mu = L.g2d_reference_mu(M, model.g)
    pde = L.g2d_reference_residual_mse(M, model.g, mu) + 1 / (M.mean(F_U, F_U) + 1e-2) + 1 / (mu ** 2 + 1e-6)
    assert abs(float(mu) - float(g["f64.mu"])) < 1e-10 * abs(float(g["f64.mu"]))
    assert abs(float(pde) - float(g["f64.pde"])) < 1e-10 * abs(float(g["f64.pde"]))
    assert abs(float(L.riesz_energy(M, model.g, "sum")) - float(g["f64.riesz"])) < 1e-10 * abs(float(g["f64.riesz"]))
    # pointwise semantics against the oracle's intended-maths function
    Xr = X.clone().requires_grad_(True)
    _, _, pde_r, _, mu_r = po.gpe2d_pointwise_terms(model, Xr, torch.tensor(g["Xb"], dtype=torch.float64))
    mu_p = L.energy_mu(M, model.g)
    pde_p = L.residual_mse(M, L.hamiltonian_weights(-1.0, model.g), mu_p) + 1 / (M.mean(F_U, F_U) + 1e-2) + 1 / (mu_p ** 2 + 1e-6)
    assert abs(float(mu_p) - float(mu_r)) < 1e-10 * abs(float(mu_r))
    assert abs(float(pde_p) - float(pde_r)) < 1e-10 * abs(float(pde_r))


@pytest.mark.parametrize("name", ["helmholtz_leaky", "helmholtz_tanh"])
def test_helmholtz_moment_algebra(golden, name):
    from gpe_b200 import losses as L
    g = golden(name)
    model = load_weights(po.Helmholtz2D([int(v) for v in g["layers"]], activation=str(g["act"])), g).double()
    X = torch.tensor(g["X"], dtype=torch.float64)
    k = float(g["k"])
    u, gr, H = oracle_jets(model, X)
    q = k ** 2 * torch.sin(np.pi * X[:, 0]) * torch.sin(np.pi * X[:, 1])
    z = torch.zeros_like(u)
    M = L.Moments(torch_moments(u, gr[:, 0], gr[:, 1], H[:, 0, 0] + H[:, 1, 1], z, z, q))
    assert abs(float(L.helmholtz_reference_mse(M, k ** 2)) - float(g["f64.pde"])) < 1e-10 * abs(float(g["f64.pde"]))
    assert abs(float(L.helmholtz_pointwise_mse(M, k ** 2)) - float(model.loss_pde_pointwise(X, k))) < 1e-10 * abs(float(g["f64.pde"]))


@pytest.mark.parametrize("name", ["root_m0_p2", "root_m2_p3"])
def test_root_rayleigh_moment_algebra(golden, name):
    from gpe_b200 import losses as L
    from gpe_b200.models import hermite_base_jets
    g = golden(name)
    mode, p, gamma = int(g["mode"]), int(g["p"]), float(g["gamma"])
    model = load_weights(po.RootGPE1D([int(v) for v in g["layers"]], mode=mode, gamma=gamma, power=p), g).double()
    X = torch.tensor(g["X"], dtype=torch.float64)
    n0, gr, H = oracle_jets(model, X)
    base = hermite_base_jets(X, mode, const_h=True).double()
    u, ux, uxx = base[:, 0] + n0, base[:, 1] + gr[:, 0], base[:, 2] + H[:, 0, 0]
    V = 0.5 * X[:, 0] ** 2
    M = L.Moments(torch_moments(u, ux, torch.zeros_like(u), uxx, V, u ** p, torch.zeros_like(u)))
    wH = L.hamiltonian_weights(-0.5, gamma)
    mu = L.rayleigh_mu(M, wH)
    # base jets are float32-rounded: agreement to ~1e-6
    assert abs(float(mu) - float(g["f64.mu"])) < 5e-6 * abs(float(g["f64.mu"]))
    assert abs(float(L.residual_mse(M, wH, mu)) - float(g["f64.pde"])) < 5e-6 * abs(float(g["f64.pde"]))


# ---------------------------------------------------------------------------------------------
# grad_fn scale tracer
# ---------------------------------------------------------------------------------------------
class JetFunctionStandIn(torch.autograd.Function):
    """CPU stand-in whose backward node name contains 'JetFunction', like the C++ node's."""

    @staticmethod
    def forward(ctx, x):
        return x * 1.0

    @staticmethod
    def backward(ctx, g):
        return g


def test_trace_scale():
    from gpe_b200.models import trace_scale
    x = torch.randn(5, requires_grad=True)
    u = JetFunctionStandIn.apply(x).unsqueeze(1)
    assert trace_scale(u) == 1.0
    assert abs(trace_scale(0.01 * u / 0.37) - 0.01 / 0.37) < 1e-12       # R/harmonic_pinn_simulation.py:333-340
    assert trace_scale(-(u * 2.0)) == -2.0
    assert trace_scale(u + 1.0) is None and trace_scale(torch.tanh(u)) is None and trace_scale(u.detach()) is None
    f = torch.sin(torch.linspace(0, 1, 5)).unsqueeze(1)
    assert trace_scale(u * f) is None
    assert abs(trace_scale(3.0 * (u * f), allow_factor=True) - 3.0) < 1e-12   # box boundary factor


# ---------------------------------------------------------------------------------------------
# data parallelism over collocation points: world_size 2 over gloo
# ---------------------------------------------------------------------------------------------
def _dp_worker(rank, world, port, golden_path, out):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from gpe_b200 import distributed as gdist, losses as L
        g = dict(np.load(golden_path, allow_pickle=False))
        model = load_weights(po.GPE2D([int(v) for v in g["layers"]], g=float(g["g"])), g).double()
        Xall = torch.tensor(g["X"], dtype=torch.float64)
        lo, hi = gdist.shard_bounds(Xall.shape[0], rank, world)
        assert gdist.shard_points(Xall).shape[0] == hi - lo
        X = Xall[lo:hi].clone().requires_grad_(True)
        # local jets by the oracle's autograd (stands in for the kernels), local moments, collective #1
        u = model.forward(X)
        gr = po.d_dx(u, X)
        lap = po.d_dx(gr[:, [0]], X)[:, 0] + po.d_dx(gr[:, [1]], X)[:, 1]
        V = model.compute_potential(X).detach()
        uu = u[:, 0]
        m = torch_moments(uu, gr[:, 0], gr[:, 1], lap, V, uu ** 3, torch.zeros_like(uu))
        M = L.Moments(gdist.all_reduce_sum(m))
        mu = L.g2d_reference_mu(M, model.g)
        loss = L.g2d_reference_residual_mse(M, model.g, mu) + L.riesz_energy(M, model.g, "sum")
        loss.backward()
        gdist.all_reduce_gradients(model.parameters())          # collective #2
        flat = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
        if rank == 0:
            torch.save({"loss": loss.detach(), "mu": mu.detach(), "grad": flat, "n": M.n.detach()}, out)
    finally:
        dist.destroy_process_group()


def test_sharded_equals_single_process_gloo(tmp_path):
    import torch.multiprocessing as mp
    from gpe_b200 import losses as L
    golden_path = os.path.join(ROOT, "tests", "golden", "g2d_g100.npz")
    out = str(tmp_path / "dp.pt")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_dp_worker, args=(2, port, golden_path, out), nprocs=2, join=True)
    got = torch.load(out)
    g = dict(np.load(golden_path, allow_pickle=False))
    model = load_weights(po.GPE2D([int(v) for v in g["layers"]], g=float(g["g"])), g).double()
    X = torch.tensor(g["X"], dtype=torch.float64, requires_grad=True)
    pde, _, mu = model.pde_loss(X, model.forward(X))
    u = model.forward(X)
    ref = pde - 1 / (torch.mean(u ** 2) + 1e-2) - 1 / (mu ** 2 + 1e-6) + model.riesz_loss(model.forward(X), X)
    gref = torch.cat([t.reshape(-1) for t in po.param_grads(model, ref)])
    assert int(got["n"]) == X.shape[0]
    assert abs(float(got["mu"]) - float(mu)) < 1e-10 * abs(float(mu))
    assert abs(float(got["loss"]) - float(ref)) < 1e-10 * abs(float(ref))
    assert float((got["grad"] - gref).norm() / gref.norm()) < 1e-9


def test_shard_bounds_cover_ragged_sizes():
    from gpe_b200.distributed import shard_bounds
    for n in (0, 1, 7, 8, 1000003):
        for world in (1, 2, 3, 8):
            spans = [shard_bounds(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def test_loss_as_epoch_learning_rate_matches_torch_scheduler():
    """gpe_b200.training._loss_as_epoch_lr restates `CosineAnnealingWarmRestarts(T_0=200, T_mult=2, eta_min=1e-6)
    .step(total_loss)` (R/harmonic_pinn_simulation.py:347-349,391) as tensor arithmetic: compare with the scheduler
    itself, both branches (loss < T_0 and the restart branch loss >= T_0)."""
    import warnings
    from torch.optim.lr_scheduler import CosineAnnealingWarmRestarts
    from gpe_b200.training import _loss_as_epoch_lr
    base = 1e-3
    opt = torch.optim.Adam([torch.nn.Parameter(torch.zeros(1))], lr=base)
    sched = CosineAnnealingWarmRestarts(opt, T_0=200, T_mult=2, eta_min=1e-6)
    for loss in [0.0, 1e-6, 0.37, 23.18, 150.0, 199.99, 200.0, 333.3, 599.9, 600.0, 1234.5, 9.9e4]:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            sched.step(torch.tensor(loss))
        want = opt.param_groups[0]["lr"]
        got = float(_loss_as_epoch_lr(torch.tensor(loss, dtype=torch.float64), base))
        assert abs(got - want) <= 1e-9 * base + 1e-6 * abs(want), (loss, got, want)   # the scheduler works on the float32 tensor


def test_airy_base_on_device_matches_the_scipy_construction():
    """gpe_b200.models.airy_base_jets (pure torch, runs where x lives) against the reference's host construction
    (R/gravity_well_pinn_simulation.py:135-160: scipy airy + ai_zeros, np.gradient), for every tabulated mode, on the
    training grid and on the two boundary points (where psi(lb) is the sign of scipy's rounding residue at the zero)."""
    from scipy.special import ai_zeros, airy
    from gpe_b200.models import airy_ai, airy_base_jets
    xs = torch.linspace(-60.0, 40.0, 50001, dtype=torch.float64)
    ai, aip = airy_ai(xs)
    ra, rap = airy(xs.numpy())[:2]
    amp = np.where(xs.numpy() < 0, np.maximum(np.abs(xs.numpy()), 1.0) ** -0.25 / np.sqrt(np.pi), np.abs(ra))
    ampp = np.where(xs.numpy() < 0, np.maximum(np.abs(xs.numpy()), 1.0) ** 0.25 / np.sqrt(np.pi), np.abs(rap))
    assert (np.abs(ai.numpy() - ra) / amp).max() < 5e-8 and (np.abs(aip.numpy() - rap) / ampp).max() < 5e-8
    for n in range(16):
        for npts in (4000, 1200, 2):
            x = torch.tensor(np.linspace(0, 35, npts).reshape(-1, 1), dtype=torch.float32)
            got = airy_base_jets(x, n).numpy()
            xn = x.reshape(-1).numpy()
            a, ap = airy(xn + ai_zeros(n + 1)[0][n])[:2]
            dx = xn[1] - xn[0]
            nf = np.sqrt(np.sum(a ** 2) * dx)
            ref = np.stack([a / nf, ap / nf, np.gradient(ap / nf, dx)], 1).astype(np.float32)
            cols = 1 if npts == 2 else 3          # airy_solution (value only) is all the two-point case feeds
            err = np.abs(got - ref)[:, :cols].max(0) / np.abs(ref)[:, :cols].max(0)
            assert err.max() < 2e-7, (n, npts, err)
    with pytest.raises(ValueError):
        airy_base_jets(torch.zeros(4, 1), 16)


# ---------------------------------------------------------------------------------------------
# collocation generators (run where they are told to: here on the CPU)
# ---------------------------------------------------------------------------------------------
def test_latin_hypercube_hits_every_stratum_once_per_dimension():
    from gpe_b200 import collocation
    n, d, lo, hi = 4096, 2, -6.0, 6.0
    X = collocation.latin_hypercube(n, d, lo, hi, seed=11, device="cpu", dtype=torch.float64)
    assert X.shape == (n, d) and float(X.min()) >= lo and float(X.max()) <= hi
    for k in range(d):
        strata = torch.floor((X[:, k] - lo) / (hi - lo) * n).long().clamp(max=n - 1)
        assert torch.equal(torch.sort(strata).values, torch.arange(n)), k          # a permutation of the n strata
    # the two dimensions are permuted independently
    sx = torch.floor((X[:, 0] - lo) / (hi - lo) * n)
    sy = torch.floor((X[:, 1] - lo) / (hi - lo) * n)
    assert float((sx == sy).double().mean()) < 0.01
    # seed-deterministic; the per-rank streams of a data-parallel run (seed + rank) differ
    assert torch.equal(X, collocation.latin_hypercube(n, d, lo, hi, seed=11, device="cpu", dtype=torch.float64))
    Y = collocation.latin_hypercube(n, d, lo, hi, seed=12, device="cpu", dtype=torch.float64)
    assert not torch.equal(X, Y) and len(set(map(tuple, X.tolist())) & set(map(tuple, Y.tolist()))) == 0
    # per-dimension bounds
    Z = collocation.latin_hypercube(1000, 2, [0.0, -1.0], [1.0, 3.0], seed=3, device="cpu", dtype=torch.float64)
    assert 0.0 <= float(Z[:, 0].min()) and float(Z[:, 0].max()) <= 1.0 and -1.0 <= float(Z[:, 1].min()) and float(Z[:, 1].max()) <= 3.0


def test_square_boundary_lies_on_the_perimeter():
    from gpe_b200 import collocation
    lo, hi = -6.0, 6.0
    P = collocation.square_boundary(500, lo, hi, seed=5, device="cpu", dtype=torch.float64)
    assert P.shape == (500, 2)
    on_edge = (P[:, 0] == lo) | (P[:, 0] == hi) | (P[:, 1] == lo) | (P[:, 1] == hi)
    assert bool(on_edge.all()) and float(P.min()) >= lo and float(P.max()) <= hi
    for col, val in ((1, lo), (1, hi), (0, lo), (0, hi)):                          # 125 points per edge
        assert int((P[:, col] == val).sum()) >= 125
    assert torch.equal(P, collocation.square_boundary(500, lo, hi, seed=5, device="cpu", dtype=torch.float64))


def test_disc_points_follow_the_reference_distribution():
    """S/gross_pitaevskii_2D.py:277-295: r ~ U(0, R), theta ~ U(0, 2 pi) -- uniform in r, NOT in area."""
    from gpe_b200 import collocation
    c, R, n = (np.pi / 2, np.pi / 2), np.pi / 2, 200_000
    P = collocation.disc_points(n, c, R, seed=2, device="cpu", dtype=torch.float64)
    r = torch.sqrt((P[:, 0] - c[0]) ** 2 + (P[:, 1] - c[1]) ** 2)
    assert float(r.max()) <= R * (1 + 1e-12)
    hist = torch.histc(r, bins=10, min=0.0, max=R) / n
    assert float((hist - 0.1).abs().max()) < 0.005                                # r uniform: an area-uniform sample would give 0.01 .. 0.19
    th = torch.atan2(P[:, 1] - c[1], P[:, 0] - c[0])
    hist_t = torch.histc(th, bins=8, min=-np.pi, max=np.pi) / n
    assert float((hist_t - 0.125).abs().max()) < 0.005
    B = collocation.disc_boundary(500, c, R, device="cpu", dtype=torch.float64)
    rb = torch.sqrt((B[:, 0] - c[0]) ** 2 + (B[:, 1] - c[1]) ** 2)
    assert float((rb - R).abs().max()) < 1e-12 and torch.allclose(B[0], B[-1], atol=1e-12)


# ---------------------------------------------------------------------------------------------
# replicated boundary term under data parallelism: counted once, not world times
# ---------------------------------------------------------------------------------------------
def _replicated_worker(rank, world, port, out):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from gpe_b200 import distributed as gdist
        torch.manual_seed(3)
        net = torch.nn.Sequential(torch.nn.Linear(2, 16), torch.nn.Tanh(), torch.nn.Linear(16, 1)).double()
        Xall = torch.rand(64, 2, dtype=torch.float64, generator=torch.Generator().manual_seed(4))
        Xb = torch.rand(10, 2, dtype=torch.float64, generator=torch.Generator().manual_seed(5))     # the same points on every rank
        lo, hi = gdist.shard_bounds(64, rank, world)
        interior = gdist.all_reduce_sum(torch.sum(net(Xall[lo:hi]) ** 2)) / 64.0                  # sharded term
        boundary = gdist.replicated_term(10.0 * torch.mean(net(Xb) ** 2))                         # replicated term
        loss = interior + boundary
        loss.backward()
        gdist.all_reduce_gradients(net.parameters())
        if rank == 0:
            torch.save({"loss": loss.detach(), "grad": torch.cat([p.grad.reshape(-1) for p in net.parameters()])}, out)
    finally:
        dist.destroy_process_group()


def test_replicated_boundary_term_is_counted_once_gloo(tmp_path):
    import torch.multiprocessing as mp
    out = str(tmp_path / "rep.pt")
    mp.spawn(_replicated_worker, args=(2, 29600 + (os.getpid() % 2000), out), nprocs=2, join=True)
    got = torch.load(out)
    torch.manual_seed(3)
    net = torch.nn.Sequential(torch.nn.Linear(2, 16), torch.nn.Tanh(), torch.nn.Linear(16, 1)).double()
    Xall = torch.rand(64, 2, dtype=torch.float64, generator=torch.Generator().manual_seed(4))
    Xb = torch.rand(10, 2, dtype=torch.float64, generator=torch.Generator().manual_seed(5))
    loss = torch.mean(net(Xall) ** 2) + 10.0 * torch.mean(net(Xb) ** 2)
    loss.backward()
    want = torch.cat([p.grad.reshape(-1) for p in net.parameters()])
    assert abs(float(got["loss"]) - float(loss)) < 1e-12 * abs(float(loss))
    assert float((got["grad"] - want).norm() / want.norm()) < 1e-12
