"""Tensor-path check on the GPU box: parity of the tcgen05 forward against the fp64 oracle + timing vs the FFMA path."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import torch
from gpe_b200 import ops
from oracle import jet_oracle as jo
from helpers import rel_err
from test_jet_kernels import make_net, points, oracle_h

print("GPE_TENSOR =", os.environ.get("GPE_TENSOR"))
for (d, W, L, act, jet, n) in [(2, 128, 4, 0, 3, 3000), (2, 100, 3, 0, 3, 1500), (1, 128, 4, 0, 1, 2049), (1, 100, 3, 0, 1, 777), (2, 100, 3, 2, 3, 800), (2, 128, 2, 0, 3, 64), (2, 128, 4, 0, 3, 100000)]:
    Ws, bs = make_net(d, W, L, seed=7 + W + L, act=act)
    x = points(n, d, seed=n)
    u, g, h = ops.mlp_jet_fwd(x.cuda(), [w.cuda() for w in Ws], [b.cuda() for b in bs], act, jet)
    torch.cuda.synchronize()
    m = min(n, 4000)
    ur, gr, Hr = jo.jets(x[:m], Ws, bs, act)
    print(f"d{d} W{W} L{L} act{act} jet{jet} n{n}: rel u {rel_err(u[:m].cpu(), ur):.2e} g {rel_err(g[:m].cpu(), gr):.2e} h {rel_err(h[:m].cpu(), oracle_h(Hr, d, jet, ops)):.2e}", flush=True)

def timed(fn, reps=5, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]

for (d, W, L, jet, n) in [(2, 128, 4, 3, 1 << 22), (1, 128, 4, 1, 1 << 22), (2, 100, 3, 3, 1 << 22)]:
    Ws, bs = make_net(d, W, L, seed=1, act=0)
    Ws = [w.cuda() for w in Ws]; bs = [b.cuda() for b in bs]
    x = torch.rand(n, d, device="cuda") * 4 - 2
    t = timed(lambda: ops.mlp_jet_fwd(x, Ws, bs, 0, jet))
    S = 1 + 2 * d
    print(f"fwd d{d} W{W} L{L} n{n}: {t:.2f} ms  {n / t / 1e3:.1f} Mpts/s  algorithmic {2 * W * W * S * (L - 1) * n / t / 1e9:.1f} TFLOP/s", flush=True)

# ---------------- backward ----------------
import numpy as np
from test_jet_kernels import hbar_full
for (d, W, L, act, jet, n) in [(2, 128, 4, 0, 3, 3000), (2, 100, 3, 0, 3, 1500), (1, 128, 4, 0, 1, 2049), (1, 100, 3, 0, 1, 777), (2, 100, 3, 2, 3, 800), (2, 128, 2, 0, 3, 64)]:
    Ws, bs = make_net(d, W, L, seed=11 + W + L, act=act)
    x = points(n, d, seed=n + 1)
    gen = torch.Generator().manual_seed(99)
    H = ops.jet_hess(d, jet)
    u_bar = torch.randn(n, generator=gen); g_bar = torch.randn((n, d), generator=gen); h_bar = torch.randn((n, H), generator=gen)
    dWs, dbs = ops.mlp_jet_bwd(x.cuda(), [w.cuda() for w in Ws], [b.cuda() for b in bs], act, jet, u_bar.cuda(), g_bar.cuda(), h_bar.cuda())
    torch.cuda.synchronize()
    rW, rb = jo.jet_functional_param_grads(x, Ws, bs, act, u_bar, g_bar, hbar_full(h_bar, d, jet, ops))
    got = np.concatenate([t.cpu().numpy().ravel() for t in dWs + dbs]); ref = np.concatenate([t.numpy().ravel() for t in rW + rb])
    per = " ".join(f"{rel_err(a.cpu(), b):.1e}" for a, b in zip(dWs + dbs, rW + rb))
    print(f"bwd d{d} W{W} L{L} act{act} n{n}: rel {rel_err(got, ref):.2e} | per tensor {per}", flush=True)

for (d, W, L, jet, n) in [(2, 128, 4, 3, 1 << 22), (1, 128, 4, 1, 1 << 22), (2, 100, 3, 3, 1 << 22)]:
    Ws, bs = make_net(d, W, L, seed=1, act=0)
    Ws = [w.cuda() for w in Ws]; bs = [b.cuda() for b in bs]
    x = torch.rand(n, d, device="cuda") * 4 - 2
    ub = torch.randn(n, device="cuda"); gb = torch.randn(n, d, device="cuda"); hb = torch.randn(n, 1, device="cuda")
    t = timed(lambda: ops.mlp_jet_bwd(x, Ws, bs, 0, jet, ub, gb, hb), reps=3, warm=1)
    S = 1 + 2 * d
    print(f"bwd d{d} W{W} L{L} n{n}: {t:.2f} ms  {n / t / 1e3:.1f} Mpts/s  algorithmic {2 * 2 * W * W * S * (L - 1) * n / t / 1e9:.1f} TFLOP/s", flush=True)
