"""Fused backward (one persistent launch, zbar through an L2-resident ring) against the per-chunk kernels: same stash, same
adjoints, gradients compared tensor by tensor; then timings.  usage: fused_check.py [n_points] [reps]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import torch
from gpe_b200 import ops
from test_jet_kernels import make_net

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
out = {"n": n}
for (d, W, L, jet) in [(2, 128, 4, 3), (1, 128, 4, 1), (2, 100, 3, 3)]:
    Ws, bs = make_net(d, W, L, seed=1, act=0)
    Ws = [w.cuda() for w in Ws]; bs = [b.cuda() for b in bs]
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.rand(n + 3, d, device="cuda", generator=g) * 4 - 2          # ragged: not a multiple of 8
    ub = torch.randn(n + 3, device="cuda", generator=g); gb = torch.randn(n + 3, d, device="cuda", generator=g)
    hb = torch.randn(n + 3, 1, device="cuda", generator=g)
    u, gg, h, stash = ops.mlp_jet_fwd_stash(x, Ws, bs, 0, jet)
    res = {}
    for mode in ("0", "1"):
        os.environ["GPE_BWD_FUSED"] = mode
        dW, db = ops.mlp_jet_bwd_stashed(x, Ws, bs, 0, jet, ub, gb, hb, stash)
        torch.cuda.synchronize()
        res[mode] = [t.double() for t in dW + db]
    errs = [float((a - b).norm() / b.norm()) for a, b in zip(res["1"], res["0"])]
    finite = all(bool(torch.isfinite(t).all()) for t in res["1"])
    key = f"d{d}_W{W}_L{L}"
    out[key] = {"fused_vs_chunked_rel_per_tensor": errs, "finite": finite}
    for mode in ("0", "1"):
        os.environ["GPE_BWD_FUSED"] = mode
        for _ in range(2):
            ops.mlp_jet_bwd_stashed(x, Ws, bs, 0, jet, ub, gb, hb, stash)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record()
        for _ in range(reps):
            ops.mlp_jet_bwd_stashed(x, Ws, bs, 0, jet, ub, gb, hb, stash)
        b.record(); torch.cuda.synchronize()
        out[key]["bwd_ms_fused" if mode == "1" else "bwd_ms_chunked"] = a.elapsed_time(b) / reps
    print(json.dumps({key: out[key]}), flush=True)
    del stash
os.environ["GPE_BWD_FUSED"] = "1"
