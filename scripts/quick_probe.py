"""First-contact probe on the GPU box: FFMA peak, fwd/bwd kernel timings at the BASELINE shapes."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import torch
from gpe_b200 import ops


def timed(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


def net(d, W, L):
    torch.manual_seed(0)
    widths = [d] + [W] * L + [1]
    Ws = [torch.randn(widths[i + 1], widths[i], device="cuda") * (1.0 / widths[i]) ** 0.5 for i in range(len(widths) - 1)]
    bs = [torch.zeros(widths[i + 1], device="cuda") + 0.01 for i in range(len(widths) - 1)]
    return Ws, bs


out = {"gpu": torch.cuda.get_device_name(0)}
for mode in (0, 1):
    iters = 20000
    ms = timed(lambda: ops.ffma_probe(mode, iters))
    fl = ops.ffma_probe(mode, iters)
    out[f"ffma_probe_mode{mode}_tflops"] = fl / ms / 1e9
for (d, W, L, jet, n) in [(1, 128, 4, 1, 1 << 20), (2, 128, 4, 3, 1 << 20), (2, 100, 3, 3, 1 << 20), (1, 100, 3, 1, 10000), (1, 128, 4, 0, 1 << 20)]:
    Ws, bs = net(d, W, L)
    x = torch.rand(n, d, device="cuda") * 4 - 2
    H = ops.jet_hess(d, jet)
    ub = torch.randn(n, device="cuda"); gb = torch.randn(n, d, device="cuda") if jet else None
    hb = torch.randn(n, H, device="cuda") if H else None
    tf = timed(lambda: ops.mlp_jet_fwd(x, Ws, bs, 0, jet))
    tb = timed(lambda: ops.mlp_jet_bwd(x, Ws, bs, 0, jet, ub, gb, hb))
    S = {0: 1, 1: 1 + 2 * d, 2: 6, 3: 2 + d}[jet]
    ffwd = 2 * W * W * S * (L - 1)
    out[f"d{d}_W{W}_L{L}_jet{jet}_n{n}"] = dict(fwd_ms=tf, bwd_ms=tb, fwd_tflops=ffwd * n / tf / 1e9,
                                                bwd_tflops=3 * ffwd * n / tb / 1e9, mpts_s_step=n / (tf + tb) / 1e3)
print(json.dumps(out, indent=1))
