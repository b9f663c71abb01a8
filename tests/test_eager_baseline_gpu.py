"""What users get today: the reference's own PyTorch code for BASELINE config 4 run with device='cuda'
on the same B200, beside the fused step, at a size the eager autograd graph fits (its ~40 [N,128] tensors per layer).
The oracle restatement *is* that code (bit-equal to the real class in float32, tests/test_oracle_golden.py), moved to
the GPU unchanged.  Writes gpurun_out/eager_vs_fused.json; asserts agreement of the two losses and that the fused step
is the faster one."""
import json
import os

import pytest
import torch

from oracle import pinn_oracle as po

pytestmark = pytest.mark.gpu


def _time_ms(fn, warm=3, reps=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(reps):
        fn()
    t1.record()
    torch.cuda.synchronize()
    return t0.elapsed_time(t1) / reps


def test_eager_reference_on_gpu_vs_fused_step():
    import gpe_b200
    layers, n, nb = [2, 128, 128, 128, 128, 1], 262144, 4096
    torch.manual_seed(1234)
    ref = po.GPE2D(layers, g=100.0, potential="harmonic").apply(po.xavier_bias001_).cuda()
    fused = gpe_b200.GPE2D(layers, g=100.0, potential="harmonic", semantics="pointwise")
    fused.load_state_dict(ref.state_dict())
    fused = fused.cuda()
    X = (torch.rand(n, 2, device="cuda") * 12 - 6).requires_grad_(True)
    Xb = torch.rand(nb, 2, device="cuda") * 12 - 6
    out = {}

    def eager_step():
        ref.zero_grad(set_to_none=True)
        total = po.gpe2d_pointwise_terms(ref, X, Xb)[0]
        total.backward()
        out["eager"] = total.detach()

    def fused_step():
        fused.zero_grad(set_to_none=True)
        total = fused.loss(X, Xb, None)
        total.backward()
        out["fused"] = total.detach()

    torch.cuda.reset_peak_memory_stats()
    t_eager = _time_ms(eager_step)
    mem_eager = torch.cuda.max_memory_allocated()
    torch.cuda.reset_peak_memory_stats()
    t_fused = _time_ms(fused_step)
    mem_fused = torch.cuda.max_memory_allocated()
    le, lf = float(out["eager"]), float(out["fused"])
    assert abs(le - lf) <= 1e-4 * abs(le), (le, lf)          # two float32 evaluations of the same loss
    ge = torch.cat([p.grad.reshape(-1) for p in ref.parameters()]).double()
    gf = torch.cat([p.grad.reshape(-1) for p in fused.parameters()]).double()
    assert float((ge - gf).norm() / ge.norm()) < 1e-3       # eager float32 autograd at N = 262144 is the noisy side
    rec = {"workload": "config 4 network, pointwise terms, N=%d (+%d boundary points)" % (n, nb),
           "eager_reference_cuda_ms": t_eager, "fused_ms": t_fused,
           "eager_points_per_s": n / t_eager * 1e3, "fused_points_per_s": n / t_fused * 1e3,
           "speedup": t_eager / t_fused, "eager_peak_bytes": mem_eager, "fused_peak_bytes": mem_fused,
           "gpu": torch.cuda.get_device_name(0)}
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/eager_vs_fused.json", "w") as fh:
        json.dump(rec, fh, indent=1)
    print(json.dumps(rec))
    assert t_fused < t_eager
