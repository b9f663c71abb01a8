"""Kernel-name histogram of one train step of a BASELINE config (CUPTI records via torch.profiler).  usage: launch_hist.py [config]"""
import collections, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import torch
import gpe_b200
import bench
cfg = int(sys.argv[1]) if len(sys.argv) > 1 else 1
wl = bench.Workload(cfg)
n = min(wl.n, 1 << 20)
model, loss_fn, X = wl.make(gpe_b200, n, torch.device("cuda"), 1234, gpu=True)
params = list(model.parameters())
def step():
    for p in params: p.grad = None
    loss_fn(X).backward()
for _ in range(3): step()
from torch.profiler import ProfilerActivity, profile
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    step(); torch.cuda.synchronize()
c = collections.Counter(); t = collections.Counter()
for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CUDA:
        name = ev.name[:90]; c[name] += 1; t[name] += ev.device_time
print("config", cfg, "n", n, "device activities", sum(c.values()))
for k, v in c.most_common(40): print(f"{v:4d}  {t[k]:9.1f} us  {k}")
