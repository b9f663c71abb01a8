"""Kernel table of ONE CUDA-graph replay of the paper's curriculum epoch (4000 points, [1,64,64,64,1],
gpe_b200.training._Stage.body: loss, backward, clip, Adam, early-stop bookkeeping), from torch.profiler (CUPTI)."""
import collections, os, re, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import numpy as np, torch
from gpe_b200 import training, FinalGPE1D
from gpe_b200.graphs import GraphedStep

lb, ub, n, layers = -10.0, 10.0, int(sys.argv[1]) if len(sys.argv) > 1 else 4000, [1, 64, 64, 64, 1]
Xn = np.linspace(lb, ub, n).reshape(-1, 1)
torch.manual_seed(0)
model = FinalGPE1D(layers, mode=0, gamma=0.0).cuda()
model.apply(lambda m: training.advanced_initialization(m, 0))
X = torch.tensor(Xn, dtype=torch.float32, device="cuda").requires_grad_(True)
bp = torch.tensor([[lb], [ub]], dtype=torch.float32, device="cuda"); bv = torch.zeros((2, 1), device="cuda")
st = training._Stage(model, X, bp, bv, float(Xn[1, 0] - Xn[0, 0]), 0.0, 3, "harmonic", 0.01, 1e-5, 100000, 1e-12)
st.body(first_of_mode=True)
gs = GraphedStep(st.body, st.opt, warmup=2)
for _ in range(5): gs()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(100): gs()
b.record(); torch.cuda.synchronize()
print(f"replay {a.elapsed_time(b) / 100:.4f} ms per epoch, {n} points")
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
    gs(); torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
agg = collections.OrderedDict()
for e in evs:
    c = agg.setdefault(e.name[:90], [0, 0.0]); c[0] += 1; c[1] += e.device_time
t0 = min(e.time_range.start for e in evs); t1 = max(e.time_range.end for e in evs)
print(f"{len(evs)} device activities, span {(t1 - t0):.1f} us, sum of kernel times {sum(v[1] for v in agg.values()):.1f} us")
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:24]:
    print(f"{t:8.1f} us x{c:3d}  {k}")


def short(nm):
    m = re.search(r"(jet_\w+|dw_tc_kernel|gram_\w+|scalar_program_kernel|pack_weights_kernel|finish_grads_kernel|multi_tensor_apply_kernel|reduce_kernel|CatArray\w+|memcpy\w*|Memset|gemmk1_kernel|lpnorm_cleanup|index_\w+)", nm)
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
