"""GPU parity of the jet kernels (through the C ABI) against the nested-autograd fp64 oracle.

Tolerance: the kernels compute in fp32; the oracle in fp64.  North-star bar: 1e-5 relative
(L2 over the point set) for values/derivatives and for parameter gradients."""
import numpy as np
import pytest
import torch

from oracle import jet_oracle as jo
from helpers import rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-5


def make_net(d, W, L, seed, act, gain=1.0):
    g = torch.Generator().manual_seed(seed)
    widths = [d] + [W] * L + [1]
    Ws, bs = [], []
    for i in range(len(widths) - 1):
        fan_in, fan_out = widths[i], widths[i + 1]
        bound = gain * (6.0 / (fan_in + fan_out)) ** 0.5
        Ws.append((torch.rand((fan_out, fan_in), generator=g) * 2 - 1) * bound)
        bs.append((torch.rand((fan_out,), generator=g) * 2 - 1) * 0.1)
    return Ws, bs


def points(n, d, seed, lo=-3.0, hi=3.0):
    g = torch.Generator().manual_seed(seed)
    return torch.rand((n, d), generator=g) * (hi - lo) + lo


def oracle_h(H, d, jet, ops):
    if jet == ops.JET_DIAG:
        return torch.stack([H[:, i, i] for i in range(d)], dim=1)
    if jet == ops.JET_LAPLACIAN:
        return sum(H[:, i, i] for i in range(d)).unsqueeze(1)
    if jet == ops.JET_FULL:
        return torch.stack([H[:, 0, 0], H[:, 1, 1], H[:, 0, 1]], dim=1) if d == 2 else H[:, 0, 0].unsqueeze(1)
    return H[:, :0, 0]


def hbar_full(h_bar, d, jet, ops):
    """kernel-layout adjoint [N,H] -> oracle [N,d,d] weights on the full Hessian."""
    n = h_bar.shape[0]
    out = torch.zeros((n, d, d), dtype=torch.float64)
    if jet == ops.JET_DIAG:
        for i in range(d):
            out[:, i, i] = h_bar[:, i]
    elif jet == ops.JET_LAPLACIAN:
        for i in range(d):
            out[:, i, i] = h_bar[:, 0]
    elif jet == ops.JET_FULL:
        if d == 2:
            out[:, 0, 0], out[:, 1, 1], out[:, 0, 1] = h_bar[:, 0], h_bar[:, 1], h_bar[:, 2]
        else:
            out[:, 0, 0] = h_bar[:, 0]
    return out


CASES = [
    # d, W, L, act, jet, n
    (1, 64, 3, 0, 1, 1000),
    (1, 100, 3, 0, 1, 777),
    (1, 128, 4, 0, 1, 2049),
    (1, 64, 3, 1, 1, 4000),
    (1, 128, 4, 0, 0, 513),
    (2, 100, 3, 0, 3, 1500),
    (2, 128, 4, 0, 3, 3000),
    (2, 128, 4, 0, 1, 640),
    (2, 128, 2, 0, 2, 999),
    (2, 100, 3, 2, 3, 800),
    (2, 64, 1, 0, 3, 100),
    (2, 200, 3, 0, 3, 700),     # the reference's own Helmholtz default [2,200,200,200,1] (S/helmholtz_2D.py:378): WP = 256
    (2, 200, 3, 2, 2, 333),     # ... with its LeakyReLU and the full Hessian
    (1, 256, 2, 0, 1, 129),
    (2, 50, 5, 0, 0, 31),
    (1, 8, 2, 0, 1, 1),
    # residual-block nets of R/box_to_gaussian_pinn_simulation.py (act 3; L = 1 + 2 * blocks)
    (1, 64, 5, 3, 1, 4000),     # its own main: layers [1,64,64,64,1] -> two blocks
    (1, 64, 3, 3, 1, 333),
    (1, 100, 7, 3, 1, 1025),
    (1, 128, 5, 3, 0, 2),       # value-only (the two boundary points)
    (1, 32, 1, 3, 1, 65),       # no block at all: layers [1,32,1]
]


@pytest.mark.parametrize("d,W,L,act,jet,n", CASES)
def test_jet_fwd(d, W, L, act, jet, n):
    from gpe_b200 import ops
    Ws, bs = make_net(d, W, L, seed=7 + W + L, act=act)
    x = points(n, d, seed=n)
    u, g, h = ops.mlp_jet_fwd(x.cuda(), [w.cuda() for w in Ws], [b.cuda() for b in bs], act, jet)
    torch.cuda.synchronize()
    ur, gr, Hr = jo.jets(x, Ws, bs, act)
    assert rel_err(u.cpu(), ur) < TOL
    if jet != ops.JET_VALUE:
        assert rel_err(g.cpu(), gr) < TOL
        assert rel_err(h.cpu(), oracle_h(Hr, d, jet, ops)) < TOL


@pytest.mark.parametrize("d,W,L,act,jet,n", CASES)
def test_jet_bwd(d, W, L, act, jet, n):
    from gpe_b200 import ops
    Ws, bs = make_net(d, W, L, seed=11 + W + L, act=act)
    x = points(n, d, seed=n + 1)
    gen = torch.Generator().manual_seed(99)
    H = ops.jet_hess(d, jet)
    u_bar = torch.randn(n, generator=gen)
    g_bar = torch.randn((n, d), generator=gen) if jet != ops.JET_VALUE else None
    h_bar = torch.randn((n, H), generator=gen) if H else None
    dWs, dbs = ops.mlp_jet_bwd(x.cuda(), [w.cuda() for w in Ws], [b.cuda() for b in bs], act, jet, u_bar.cuda(),
                               None if g_bar is None else g_bar.cuda(), None if h_bar is None else h_bar.cuda())
    torch.cuda.synchronize()
    gb = torch.zeros((n, d)) if g_bar is None else g_bar
    hb = torch.zeros((n, d, d), dtype=torch.float64) if h_bar is None else hbar_full(h_bar, d, jet, ops)
    rW, rb = jo.jet_functional_param_grads(x, Ws, bs, act, u_bar, gb, hb)
    got = np.concatenate([t.cpu().numpy().ravel() for t in dWs + dbs])
    ref = np.concatenate([t.numpy().ravel() for t in rW + rb])
    assert rel_err(got, ref) < TOL
    for a, b in zip(dWs + dbs, rW + rb):   # per-tensor too (looser: small tensors, fp32 sums over n)
        if float(b.norm()) > 1e-6:
            assert rel_err(a.cpu(), b) < 5 * TOL


def test_mirror_xform():
    from gpe_b200 import ops
    Ws, bs = make_net(1, 64, 3, seed=3, act=0)
    x = points(500, 1, seed=5)
    cW, cb = [w.cuda() for w in Ws], [b.cuda() for b in bs]
    u, _, _ = ops.mlp_jet_fwd(x.cuda(), cW, cb, 0, ops.JET_VALUE, xform=([-1.0], [0.5]))
    ur, _, _ = jo.jets(0.5 - x, Ws, bs, 0)
    assert rel_err(u.cpu(), ur) < TOL


def test_empty_and_errors():
    from gpe_b200 import ops
    Ws, bs = make_net(1, 64, 3, seed=3, act=0)
    cW, cb = [w.cuda() for w in Ws], [b.cuda() for b in bs]
    u, g, h = ops.mlp_jet_fwd(torch.empty((0, 1), device="cuda"), cW, cb, 0, ops.JET_DIAG)
    assert u.shape == (0,) and g.shape == (0, 1)
    with pytest.raises(RuntimeError):
        ops.mlp_jet_fwd(torch.zeros((4, 1)), cW, cb, 0, ops.JET_DIAG)      # CPU tensor: no fallback
    with pytest.raises((ValueError, RuntimeError)):
        ops.mlp_jet_fwd(torch.zeros((4, 2), device="cuda"), cW, cb, 0, ops.JET_DIAG)
    with pytest.raises(RuntimeError):       # residual blocks come in (lin1, lin2) pairs: n_hidden = 1 + 2 * blocks
        Ws4, bs4 = make_net(1, 64, 4, seed=3, act=0)
        ops.mlp_jet_fwd(torch.zeros((4, 1), device="cuda"), [w.cuda() for w in Ws4], [b.cuda() for b in bs4],
                        ops.ACT_RESBLOCK_TANH, ops.JET_DIAG)
    with pytest.raises(RuntimeError):       # ... and the reference's residual network is 1-D
        Ws2, bs2 = make_net(2, 64, 3, seed=3, act=0)
        ops.mlp_jet_fwd(torch.zeros((4, 2), device="cuda"), [w.cuda() for w in Ws2], [b.cuda() for b in bs2],
                        ops.ACT_RESBLOCK_TANH, ops.JET_LAPLACIAN)


# ---------------------------------------------------------------------------------------------
# both kernel families: the tcgen05 3xTF32 path (default where eligible) and the FFMA path must each
# meet the oracle bar on the shapes the tensor path accepts, and agree with each other
# ---------------------------------------------------------------------------------------------
TC_CASES = [c for c in CASES if c[1] > 64 and c[2] - 1 <= 3 and c[2] >= 2 and (c[4] == 3 or (c[0] == 1 and c[4] in (1, 2)))]
TC_CASES += [(2, 128, 4, 1, 3, 4097), (2, 96, 2, 0, 3, 65), (1, 128, 3, 0, 1, 262144 + 77)]   # shifted tanh; ragged; two chunks


@pytest.mark.parametrize("d,W,L,act,jet,n", TC_CASES)
def test_tensor_and_ffma_paths(d, W, L, act, jet, n):
    from gpe_b200 import ops
    Ws, bs = make_net(d, W, L, seed=21 + W + L, act=act)
    x = points(n, d, seed=n + 3)
    m = min(n, 3000)                        # oracle (nested autograd, fp64) on a prefix; path-vs-path on everything
    gen = torch.Generator().manual_seed(5)
    u_bar, g_bar, h_bar = torch.randn(n, generator=gen), torch.randn((n, d), generator=gen), torch.randn((n, 1), generator=gen)
    if n > m:                               # adjoints vanish outside the prefix, so the oracle gradient is the full one
        u_bar[m:], g_bar[m:], h_bar[m:] = 0, 0, 0
    cW, cb = [w.cuda() for w in Ws], [b.cuda() for b in bs]
    out = {}
    prev = ops.set_tensor_path(True)
    try:
        for name, flag in (("tensor", True), ("ffma", False)):
            ops.set_tensor_path(flag)
            u, g, h = ops.mlp_jet_fwd(x.cuda(), cW, cb, act, jet)
            dWs, dbs = ops.mlp_jet_bwd(x.cuda(), cW, cb, act, jet, u_bar.cuda(), g_bar.cuda(), h_bar.cuda())
            torch.cuda.synchronize()
            out[name] = (u.cpu(), g.cpu(), h.cpu(), np.concatenate([t.cpu().numpy().ravel() for t in dWs + dbs]))
    finally:
        ops.set_tensor_path(prev)
    ur, gr, Hr = jo.jets(x[:m], Ws, bs, act)
    rW, rb = jo.jet_functional_param_grads(x[:m], Ws, bs, act, u_bar[:m], g_bar[:m], hbar_full(h_bar[:m], d, ops.JET_LAPLACIAN, ops))
    ref = np.concatenate([t.numpy().ravel() for t in rW + rb])
    for name in ("tensor", "ffma"):
        u, g, h, grads = out[name]
        assert rel_err(u[:m], ur) < TOL, name
        assert rel_err(g[:m], gr) < TOL, name
        assert rel_err(h[:m], oracle_h(Hr, d, ops.JET_LAPLACIAN, ops)) < TOL, name
        assert rel_err(grads, ref) < TOL, name
    for a, b in zip(out["tensor"][:3], out["ffma"][:3]):
        assert rel_err(a, b) < TOL


@pytest.mark.parametrize("d,W,L,n", [(2, 128, 4, 70001), (1, 100, 3, 4099), (2, 128, 2, 13)])
def test_resident_stash_matches_recompute(d, W, L, n):
    """gpe_mlp_jet_fwd_stash + gpe_mlp_jet_bwd_stashed (memory) against gpe_mlp_jet_fwd + gpe_mlp_jet_bwd (recompute):
    the same kernels on the same numbers -- jets bit-identical, gradients equal up to the order of the fp32 atomics."""
    from gpe_b200 import ops
    Ws, bs = make_net(d, W, L, seed=3, act=ops.ACT_TANH)
    Ws, bs = [w.cuda() for w in Ws], [b.cuda() for b in bs]
    x = points(n, d, seed=5).cuda()
    jet = ops.JET_LAPLACIAN
    u0, g0, h0 = ops.mlp_jet_fwd(x, Ws, bs, ops.ACT_TANH, jet)
    u1, g1, h1, stash = ops.mlp_jet_fwd_stash(x, Ws, bs, ops.ACT_TANH, jet)
    assert torch.equal(u0, u1) and torch.equal(g0, g1) and torch.equal(h0, h1)
    assert stash.dtype == torch.uint8 and stash.numel() == (L - 1) * ((n + 7) // 8) * 16384
    torch.manual_seed(1)
    ub, gb, hb = torch.randn(n, device="cuda"), torch.randn(n, d, device="cuda"), torch.randn(n, 1, device="cuda")
    dW0, db0 = ops.mlp_jet_bwd(x, Ws, bs, ops.ACT_TANH, jet, ub, gb, hb)
    dW1, db1 = ops.mlp_jet_bwd_stashed(x, Ws, bs, ops.ACT_TANH, jet, ub, gb, hb, stash)
    for a, b in zip(dW0 + db0, dW1 + db1):   # (the scalar output-bias gradient is a cancelling sum of 4099 atomics: ~2e-6)
        assert rel_err(a.cpu(), b.cpu()) < 1e-5
    with pytest.raises(RuntimeError):   # a stash of the wrong size is refused, not read out of bounds
        ops.mlp_jet_bwd_stashed(x, Ws, bs, ops.ACT_TANH, jet, ub, gb, hb, stash[:-16384].contiguous())
