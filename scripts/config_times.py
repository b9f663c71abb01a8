"""Train-step time of the other BASELINE shapes through the module API (not bench lines: parity-test cases whose
throughput is reported in DESIGN.md).  config 2: 1-D F-opt total loss, [1,128x4,1], 1M points; config 3: Helmholtz
[2,100x3,1], 4M points + 400 boundary points."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import numpy as np, torch
import gpe_b200
from gpe_b200 import collocation

def timed(step, reps=5, warm=3):
    for _ in range(warm): step()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): step()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps

def xavier(m):
    torch.manual_seed(1234)
    for mod in m.modules():
        if isinstance(mod, torch.nn.Linear):
            torch.nn.init.xavier_uniform_(mod.weight); mod.bias.data.fill_(0.01)
    return m

out = {}
# config 3: Helmholtz
n = 4_000_000
model = xavier(gpe_b200.Helmholtz2D([2, 100, 100, 100, 1], activation="tanh", semantics="pointwise")).cuda()
X = collocation.latin_hypercube(n, 2, 0.0, float(np.pi), seed=1234).requires_grad_(True)
Xu = collocation.square_boundary(400, 0.0, float(np.pi), seed=1)
Uu = torch.zeros(400, 1, device="cuda")
k = float(np.sqrt(2.0) * np.pi)
def step3():
    model.zero_grad(set_to_none=True)
    loss = model.compute_loss(X, Xu, Uu, k); loss.backward()
ms = timed(step3); out["config3_helmholtz_4M"] = {"ms_per_step": ms, "points_per_s": n / ms * 1e3}
# config 2: 1-D, F-opt total loss with symmetry (mode 1)
n = 1_000_000
model = xavier(gpe_b200.OptGPE1D([1, 128, 128, 128, 128, 1], g=50.0, mode=1)).cuda()
lb, ub = -10.0, 10.0
X = collocation.latin_hypercube(n, 1, lb, ub, seed=1234).requires_grad_(True)
bp = torch.tensor([[lb], [ub]], device="cuda"); bv = torch.zeros(2, 1, device="cuda")
w = [1.0, 1.0, 1.0, 1.0, 1.0]
def step2():
    model.zero_grad(set_to_none=True)
    res = model.total_loss(X, bp, bv, 50.0, lb, ub, w, "harmonic")
    loss = res[0] if isinstance(res, (tuple, list)) else res
    loss.backward()
try:
    ms = timed(step2); out["config2_1d_1M"] = {"ms_per_step": ms, "points_per_s": n / ms * 1e3}
except Exception as e:
    out["config2_1d_1M"] = {"error": repr(e)[:300]}
print(json.dumps(out))
