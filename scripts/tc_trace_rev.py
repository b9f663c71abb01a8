"""Debug probe: clock64 stamps of the two-tile reverse sweep (run with the -DGPE_TC_TRACE build copied over libgpe_b200.so)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
os.environ["GPE_TC_TWO_TILES_REV"] = "1"
import torch
from gpe_b200 import ops, _lib
from test_jet_kernels import make_net
n, d, W, L, jet = 148 * 32 * 8, 2, 128, 4, 3
Ws, bs = make_net(d, W, L, seed=1, act=0)
Ws = [w.cuda() for w in Ws]; bs = [b.cuda() for b in bs]
x = torch.rand(n, d, device="cuda") * 4 - 2
ub = torch.randn(n, device="cuda"); gb = torch.randn(n, d, device="cuda"); hb = torch.randn(n, 1, device="cuda")
u, g, h, stash = ops.mlp_jet_fwd_stash(x, Ws, bs, 0, jet)
for _ in range(3):
    ops.mlp_jet_bwd_stashed(x, Ws, bs, 0, jet, ub, gb, hb, stash)
    torch.cuda.synchronize()
buf = (ctypes.c_longlong * 4096)()
_lib.lib.gpe_debug_trace(buf, 4096)
t0 = min(v for v in buf if v)
print("chunk  mma_woke  mma_committed  epi_got_slot  epi_arrived")
for i in range(48):
    print(i, *[buf[4 * i + k] - t0 for k in range(4)])
