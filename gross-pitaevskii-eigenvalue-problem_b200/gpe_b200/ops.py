"""Torch custom ops (`torch.ops.gpe.*`) over the C ABI of libgpe_b200.so.

CUDA-only, fp32, contiguous tensors, launched on the current torch stream.  The ops own no
state; the device workspace comes from torch's caching allocator per call.

  gpe::mlp_jet_fwd  -- value / gradient / second derivatives of the MLP per collocation point
                       (replaces nn.Sequential forward + nested torch.autograd.grad:
                        NBroot cell 6:50-54, 90-104; S/gross_pitaevskii_2D.py:67-81, 183-188;
                        S/helmholtz_2D.py:40-56, 75-83)
  gpe::mlp_jet_bwd  -- parameter gradients of a linear functional of those jets
                       (replaces loss.backward(): NBroot cell 10:100, S/gross_pitaevskii_2D.py:376,422)
"""
import ctypes
from typing import List, Optional, Sequence

import torch

from . import _lib
from ._lib import (ACT_LEAKY_RELU, ACT_SHIFTED_TANH, ACT_TANH, JET_DIAG, JET_FULL, JET_LAPLACIAN,  # noqa: F401
                   JET_VALUE)

PREC_FP32, PREC_TF32X3, PREC_AUTO = 0, 1, 2


def jet_hess(d: int, jet: int) -> int:
    """Number of second-derivative outputs per point for a jet kind."""
    if jet == JET_VALUE:
        return 0
    if jet == JET_DIAG:
        return d
    if jet == JET_FULL:
        return 3 if d == 2 else 1
    return 1


def _desc(Ws, act, jet, precision):
    d = Ws[0].shape[1]
    L = len(Ws) - 1
    width = Ws[0].shape[0]
    for i, W in enumerate(Ws):
        exp = (width, d) if i == 0 else ((1, width) if i == L else (width, width))
        if tuple(W.shape) != exp:
            raise ValueError(f"layer {i}: weight shape {tuple(W.shape)} != {exp} (uniform hidden width required)")
    return _lib.MlpDesc(d, L, width, act, jet, precision)


def _check(t, name, shape=None):
    if t is None:
        return
    if not t.is_cuda:
        raise RuntimeError(f"gpe_b200: `{name}` must be a CUDA tensor (there is no CPU path)")
    if t.dtype != torch.float32:
        raise TypeError(f"gpe_b200: `{name}` must be float32, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"gpe_b200: `{name}` must be contiguous")
    if shape is not None and tuple(t.shape) != tuple(shape):
        raise ValueError(f"gpe_b200: `{name}` has shape {tuple(t.shape)}, expected {tuple(shape)}")


def _ptr(t):
    return None if t is None else t.data_ptr()


def _xform(xform):
    if xform is None:
        return None
    a, b = xform
    xf = _lib.XForm()
    for i in range(2):
        xf.a[i] = float(a[i]) if i < len(a) else 1.0
        xf.b[i] = float(b[i]) if i < len(b) else 0.0
    return ctypes.byref(xf)


def _workspace(desc, n, device, which):
    fwd, bwd = ctypes.c_size_t(0), ctypes.c_size_t(0)
    _lib.check(_lib.lib.gpe_workspace_bytes(ctypes.byref(desc), n, ctypes.byref(fwd), ctypes.byref(bwd)), "gpe_workspace_bytes")
    nbytes = fwd.value if which == "fwd" else bwd.value
    return torch.empty(max(nbytes, 16), dtype=torch.uint8, device=device), nbytes


def mlp_jet_fwd(x: torch.Tensor, Ws: Sequence[torch.Tensor], bs: Sequence[torch.Tensor], act: int, jet: int,
                xform=None, precision: int = PREC_AUTO):
    """u [N], g [N,d], h [N,H] of the MLP at x [N,d].  `xform=(a, b)` evaluates at a*x+b
    (JET_VALUE only; the mirror points of symmetry_loss)."""
    _check(x, "x")
    if x.dim() != 2:
        raise ValueError("x must be [N,d]")
    n, d = x.shape
    desc = _desc(Ws, act, jet, precision)
    if desc.d != d:
        raise ValueError(f"x has {d} columns, the network expects {desc.d}")
    for i, (W, b) in enumerate(zip(Ws, bs)):
        _check(W, f"W[{i}]")
        _check(b, f"b[{i}]", (W.shape[0],))
    H = jet_hess(d, jet)
    u = torch.empty(n, dtype=torch.float32, device=x.device)
    g = torch.empty((n, d if jet != JET_VALUE else 0), dtype=torch.float32, device=x.device)
    h = torch.empty((n, H), dtype=torch.float32, device=x.device)
    if n == 0:
        return u, g, h
    with torch.cuda.device(x.device):
        ws, nbytes = _workspace(desc, n, x.device, "fwd")
        Wp, _k1 = _lib.ptr_array([W.data_ptr() for W in Ws])
        bp, _k2 = _lib.ptr_array([b.data_ptr() for b in bs])
        stream = torch.cuda.current_stream(x.device).cuda_stream
        rc = _lib.lib.gpe_mlp_jet_fwd(ctypes.byref(desc), Wp, bp, x.data_ptr(), n, _xform(xform), u.data_ptr(),
                                      _ptr(g) if jet != JET_VALUE else None, _ptr(h) if H else None,
                                      ws.data_ptr(), nbytes, stream)
    _lib.check(rc, "gpe_mlp_jet_fwd")
    return u, g, h


def mlp_jet_bwd(x: torch.Tensor, Ws: Sequence[torch.Tensor], bs: Sequence[torch.Tensor], act: int, jet: int,
                u_bar: Optional[torch.Tensor], g_bar: Optional[torch.Tensor], h_bar: Optional[torch.Tensor],
                xform=None, precision: int = PREC_AUTO):
    """Gradients (dW list, db list) of  sum_n( u_bar u + <g_bar, g> + <h_bar, h> )."""
    _check(x, "x")
    n, d = x.shape
    desc = _desc(Ws, act, jet, precision)
    H = jet_hess(d, jet)
    _check(u_bar, "u_bar", (n,))
    _check(g_bar, "g_bar", (n, d))
    _check(h_bar, "h_bar", (n, H))
    for i, (W, b) in enumerate(zip(Ws, bs)):
        _check(W, f"W[{i}]")
        _check(b, f"b[{i}]", (W.shape[0],))
    dWs = [torch.empty_like(W) for W in Ws]
    dbs = [torch.empty_like(b) for b in bs]
    with torch.cuda.device(x.device):
        ws, nbytes = _workspace(desc, n, x.device, "bwd")
        Wp, _k1 = _lib.ptr_array([W.data_ptr() for W in Ws])
        bp, _k2 = _lib.ptr_array([b.data_ptr() for b in bs])
        dWp, _k3 = _lib.ptr_array([W.data_ptr() for W in dWs])
        dbp, _k4 = _lib.ptr_array([b.data_ptr() for b in dbs])
        stream = torch.cuda.current_stream(x.device).cuda_stream
        rc = _lib.lib.gpe_mlp_jet_bwd(ctypes.byref(desc), Wp, bp, x.data_ptr(), n, _xform(xform), _ptr(u_bar),
                                      _ptr(g_bar) if jet != JET_VALUE else None, _ptr(h_bar) if H else None, dWp, dbp,
                                      ws.data_ptr(), nbytes, stream)
    _lib.check(rc, "gpe_mlp_jet_bwd")
    return dWs, dbs


def ffma_probe(mode: int, iters: int, device=None):
    """Launch the FFMA peak probe; returns the number of fp32 FLOPs it executes."""
    device = torch.device("cuda") if device is None else device
    sink = torch.zeros(4, dtype=torch.float32, device=device)
    flops = ctypes.c_double(0)
    with torch.cuda.device(device):
        rc = _lib.lib.gpe_ffma_probe(mode, iters, sink.data_ptr(), ctypes.byref(flops),
                                     torch.cuda.current_stream(device).cuda_stream)
    _lib.check(rc, "gpe_ffma_probe")
    return flops.value


# ---------------------------------------------------------------------------------------------
# torch.library registration: torch.ops.gpe.mlp_jet_fwd / mlp_jet_bwd
# ---------------------------------------------------------------------------------------------
_LIB = torch.library.Library("gpe", "DEF")
_LIB.define("mlp_jet_fwd(Tensor x, Tensor[] W, Tensor[] b, int act, int jet, float[]? xform_a=None, "
            "float[]? xform_b=None, int precision=2) -> (Tensor, Tensor, Tensor)")
_LIB.define("mlp_jet_bwd(Tensor x, Tensor[] W, Tensor[] b, int act, int jet, Tensor? u_bar, Tensor? g_bar, "
            "Tensor? h_bar, float[]? xform_a=None, float[]? xform_b=None, int precision=2) -> (Tensor[], Tensor[])")


def _op_fwd(x, W, b, act, jet, xform_a=None, xform_b=None, precision=2):
    xf = None if xform_a is None else (xform_a, xform_b if xform_b is not None else [0.0, 0.0])
    return mlp_jet_fwd(x, list(W), list(b), act, jet, xf, precision)


def _op_bwd(x, W, b, act, jet, u_bar, g_bar, h_bar, xform_a=None, xform_b=None, precision=2):
    xf = None if xform_a is None else (xform_a, xform_b if xform_b is not None else [0.0, 0.0])
    return mlp_jet_bwd(x, list(W), list(b), act, jet, u_bar, g_bar, h_bar, xf, precision)


def _no_cpu(*a, **k):
    raise RuntimeError("gpe_b200 ops are CUDA-only: there is no CPU fallback")


_LIB.impl("mlp_jet_fwd", _op_fwd, "CUDA")
_LIB.impl("mlp_jet_bwd", _op_bwd, "CUDA")
_LIB.impl("mlp_jet_fwd", _no_cpu, "CPU")
_LIB.impl("mlp_jet_bwd", _no_cpu, "CPU")
