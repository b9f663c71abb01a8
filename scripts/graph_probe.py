"""Config 1 (1-D GPE, [1,100,100,100,1], 10k points, F-root loss, Adam + clip): eager step vs CUDA-graph replay."""
import os, sys, json, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import torch
import gpe_b200
from oracle import pinn_oracle as po

layers, lb, ub, n, gamma, p = [1, 100, 100, 100, 1], -10.0, 10.0, 10000, 1.0, 3
torch.manual_seed(0)
model = gpe_b200.RootGPE1D(layers, mode=0, gamma=gamma).cuda()
X = torch.linspace(lb, ub, n, device="cuda").view(-1, 1).requires_grad_(True)
dx = (ub - lb) / (n - 1)
bp = torch.tensor([[lb], [ub]], device="cuda"); bv = torch.zeros(2, 1, device="cuda")
opt = torch.optim.Adam(model.parameters(), lr=1e-3, capturable=True)

def step():
    opt.zero_grad(set_to_none=False)
    total, pde, bc, nrm, sym, mu = po.root_train_terms(model, X, bp, bv, dx, gamma, p, lb, ub)
    total.backward()
    torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm=1.0)
    opt.step()
    return total, mu

def wall(fn, reps):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps * 1e3

s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(5): step()
torch.cuda.current_stream().wait_stream(s)
eager_ms = wall(step, 50)
out = {"eager_ms": eager_ms, "eager_points_per_s": n / eager_ms * 1e3}
try:
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        total_s, mu_s = step()
    graph_ms = wall(g.replay, 200)
    out.update({"graph_ms": graph_ms, "graph_points_per_s": n / graph_ms * 1e3, "loss": float(total_s), "mu": float(mu_s)})
except Exception as e:
    out["graph_error"] = repr(e)[:500]
print(json.dumps(out))
