"""gpe_mlp_desc.precision is honoured per call (ABI hygiene): FP32 -> FFMA kernels, TF32X3 -> tcgen05
kernels or a loud error, AUTO -> the default rule."""
import pytest
import torch

from test_jet_kernels import make_net

pytestmark = pytest.mark.gpu


def _kernels_of(fn):
    from torch.profiler import ProfilerActivity, profile
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        fn()
        torch.cuda.synchronize()
    return [ev.name for ev in prof.events() if ev.device_type == torch.autograd.DeviceType.CUDA]


def test_precision_selects_the_kernels_per_call():
    from gpe_b200 import ops
    Ws, bs = make_net(2, 128, 3, seed=1, act=0)
    Ws, bs = [w.cuda() for w in Ws], [b.cuda() for b in bs]
    x = torch.rand(4096, 2, device="cuda") * 2 - 1
    names = {p: _kernels_of(lambda p=p: ops.mlp_jet_fwd(x, Ws, bs, 0, ops.JET_LAPLACIAN, precision=p)) for p in (0, 1, 2)}
    assert any("jet_fwd_kernel" in k for k in names[ops.PREC_FP32]) and not any("tc_kernel" in k for k in names[ops.PREC_FP32])
    assert any("tc_kernel" in k for k in names[ops.PREC_TF32X3]) and any("tc_kernel" in k for k in names[ops.PREC_AUTO])
    a = ops.mlp_jet_fwd(x, Ws, bs, 0, ops.JET_LAPLACIAN, precision=ops.PREC_FP32)
    b = ops.mlp_jet_fwd(x, Ws, bs, 0, ops.JET_LAPLACIAN, precision=ops.PREC_TF32X3)
    for u, v in zip(a, b):
        assert float((u - v).norm() / v.norm()) < 1e-5
    # widths up to 64 (the paper's nets) are zero-padded into the 128-wide tensor kernels: AUTO picks them, and they agree with FFMA
    Ws64, bs64 = make_net(2, 64, 3, seed=1, act=0)
    Ws64, bs64 = [w.cuda() for w in Ws64], [b.cuda() for b in bs64]
    k64 = _kernels_of(lambda: ops.mlp_jet_fwd(x, Ws64, bs64, 0, ops.JET_LAPLACIAN, precision=ops.PREC_AUTO))
    assert any("tc_kernel" in k for k in k64)
    a64 = ops.mlp_jet_fwd(x, Ws64, bs64, 0, ops.JET_LAPLACIAN, precision=ops.PREC_FP32)
    b64 = ops.mlp_jet_fwd(x, Ws64, bs64, 0, ops.JET_LAPLACIAN, precision=ops.PREC_TF32X3)
    for u, v in zip(a64, b64):
        assert float((u - v).norm() / v.norm()) < 1e-5
    # no tensor-core kernel for width 200 / the full 2-D Hessian: TF32X3 must refuse, AUTO and FP32 must run
    Ws200, bs200 = make_net(2, 200, 3, seed=1, act=0)
    Ws200, bs200 = [w.cuda() for w in Ws200], [b.cuda() for b in bs200]
    with pytest.raises(RuntimeError, match="GPE_PREC_TF32X3"):
        ops.mlp_jet_fwd(x, Ws200, bs200, 0, ops.JET_LAPLACIAN, precision=ops.PREC_TF32X3)
    with pytest.raises(RuntimeError, match="GPE_PREC_TF32X3"):
        ops.mlp_jet_fwd(x, Ws, bs, 0, ops.JET_FULL, precision=ops.PREC_TF32X3)
    ops.mlp_jet_fwd(x, Ws200, bs200, 0, ops.JET_LAPLACIAN, precision=ops.PREC_AUTO)
    with pytest.raises(RuntimeError, match="unknown precision"):
        ops.mlp_jet_fwd(x, Ws, bs, 0, ops.JET_LAPLACIAN, precision=7)
