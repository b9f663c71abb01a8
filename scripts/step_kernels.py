"""Kernel table of ONE CUDA-graph replay of a small training step (BASELINE config 1: loss, backward, clip, Adam), from
torch.profiler (CUPTI) -- what is left in the launch-bound regime.  usage: step_kernels.py [config] [points]"""
import gc, os, sys, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import torch
import gpe_b200
from gpe_b200.graphs import GraphedStep
import bench

cfg = int(sys.argv[1]) if len(sys.argv) > 1 else 1
wl = bench.Workload(cfg)
n = int(sys.argv[2]) if len(sys.argv) > 2 else wl.n
dev = torch.device("cuda:0")
model, loss_fn, X = wl.make(gpe_b200, n, dev, 1234, True)[:3]
params = [p for p in model.parameters() if p.requires_grad]
gopt = torch.optim.Adam(params, lr=wl.lr, capturable=True)


def body():
    gopt.zero_grad(set_to_none=True)
    loss = loss_fn(X)
    loss.backward()
    torch.nn.utils.clip_grad_norm_(params, max_norm=1.0)
    gopt.step()
    return loss


gc.collect()
gs = GraphedStep(body, gopt)
for _ in range(5): gs()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(50): gs()
b.record(); torch.cuda.synchronize()
print(f"replay {a.elapsed_time(b) / 50:.4f} ms per step, {n} points")
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
    gs(); torch.cuda.synchronize()
agg = collections.OrderedDict()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
t0 = min(e.time_range.start for e in evs); t1 = max(e.time_range.end for e in evs)
for e in evs:
    k = e.name[:90]
    c = agg.setdefault(k, [0, 0.0]); c[0] += 1; c[1] += e.device_time
print(f"{len(evs)} device activities, span {(t1 - t0):.1f} us, sum of kernel times {sum(v[1] for v in agg.values()):.1f} us")
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:30]:
    print(f"{t:8.1f} us x{c:3d}  {k}")
if os.environ.get("GPE_STEP_SEQ"):
    import re
    def short(nm):
        m = re.search(r"(jet_\w+|dw_tc_kernel|gram_\w+|scalar_program_kernel|pack_weights_kernel|finish_grads_kernel|multi_tensor_apply_kernel|reduce_kernel|CatArray\w+|memcpy\w*|Memset|gemmk1_kernel|lpnorm_cleanup)", nm)
        if m: return m.group(1)
        m = re.search(r"native::(?:\(anonymous namespace\)::)?(\w+)<?", nm.split("elementwise_kernel<")[-1])
        return "ew:" + (m.group(1) if m else nm[:30])
    seq = [short(e.name) for e in sorted(evs, key=lambda e: e.time_range.start)]
    out, i = [], 0
    while i < len(seq):
        j = i
        while j < len(seq) and seq[j] == seq[i]: j += 1
        out.append(seq[i] + (f" x{j - i}" if j - i > 1 else "")); i = j
    print(" | ".join(out))
