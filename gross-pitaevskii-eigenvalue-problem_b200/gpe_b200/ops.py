"""Torch custom ops (`torch.ops.gpe.*`) of gpe_b200.

The ops are defined in C++ (csrc/torch_shim.cpp -> libgpe_b200_torch.so, TORCH_LIBRARY(gpe)) and
forward to the C ABI of libgpe_b200.so (include/gpe_b200.h).  CUDA-only, fp32, contiguous tensors,
launched on the current torch stream; no state, workspace from torch's caching allocator per call.

  gpe::mlp_jet_fwd   value / gradient / second derivatives of the MLP per collocation point
                     (replaces nn.Sequential forward + nested torch.autograd.grad:
                      NBroot cell 6:50-54, 90-104; S/gross_pitaevskii_2D.py:67-81, 183-188;
                      S/helmholtz_2D.py:40-56, 75-83)
  gpe::mlp_jet_bwd   parameter gradients of a linear functional of those jets
                     (replaces loss.backward(): NBroot cell 10:100, S/gross_pitaevskii_2D.py:376,422)
  gpe::mlp_jets      the differentiable node built from the two (C++ autograd::Function)
  gpe::gram_moments / gpe::gram_adjoints
                     fused loss epilogue: fp64 Gram moments of the per-point features and the
                     per-point adjoints for a given dL/dm (replaces the elementwise + mean/sum part of
                     pde_loss / riesz_loss / normalization_loss / symmetry_loss / loss_pde)
"""
import os
from typing import Optional, Sequence

import torch

from . import _lib
from ._lib import (ACT_LEAKY_RELU, ACT_RESBLOCK_TANH, ACT_SHIFTED_TANH, ACT_TANH, JET_DIAG, JET_FULL, JET_LAPLACIAN,  # noqa: F401
                   JET_VALUE, NL_ABSPOW, NL_NONE, NL_POW, NMOM)

PREC_FP32, PREC_TF32X3, PREC_AUTO = 0, 1, 2

SHIM_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libgpe_b200_torch.so")
if not os.path.exists(SHIM_PATH):
    raise ImportError(f"{SHIM_PATH} is missing: build it with `make -C gross-pitaevskii-eigenvalue-problem_b200/csrc`. "
                      "gpe_b200 has no CPU / pure-PyTorch fallback.")
torch.ops.load_library(SHIM_PATH)
_ops = torch.ops.gpe


def jet_hess(d: int, jet: int) -> int:
    """Number of second-derivative outputs per point for a jet kind."""
    if jet == JET_VALUE:
        return 0
    if jet == JET_DIAG:
        return d
    if jet == JET_FULL:
        return 3 if d == 2 else 1
    return 1


def _xf(xform):
    if xform is None:
        return None, None
    a, b = xform
    return [float(v) for v in a], [float(v) for v in b]


def mlp_jet_fwd(x: torch.Tensor, Ws: Sequence[torch.Tensor], bs: Sequence[torch.Tensor], act: int, jet: int,
                xform=None, precision: int = PREC_AUTO):
    """u [N], g [N,d], h [N,H] of the MLP at x [N,d] (no autograd).  `xform=(a, b)` evaluates at
    a*x+b (JET_VALUE only; the mirror points of symmetry_loss)."""
    xa, xb = _xf(xform)
    return _ops.mlp_jet_fwd(x, list(Ws), list(bs), act, jet, xa, xb, precision)


def mlp_jet_bwd(x: torch.Tensor, Ws: Sequence[torch.Tensor], bs: Sequence[torch.Tensor], act: int, jet: int,
                u_bar: Optional[torch.Tensor], g_bar: Optional[torch.Tensor], h_bar: Optional[torch.Tensor],
                xform=None, precision: int = PREC_AUTO):
    """Gradients (dW list, db list) of  sum_n( u_bar u + <g_bar, g> + <h_bar, h> )."""
    xa, xb = _xf(xform)
    out = _ops.mlp_jet_bwd(x, list(Ws), list(bs), act, jet, u_bar, g_bar, h_bar, xa, xb, precision)
    n = len(Ws)
    return list(out[:n]), list(out[n:])


def mlp_jet_fwd_stash(x: torch.Tensor, Ws: Sequence[torch.Tensor], bs: Sequence[torch.Tensor], act: int, jet: int,
                      precision: int = PREC_AUTO):
    """(u, g, h, stash): the forward that also keeps the pre-activation jets of all points (uint8 tensor of
    gpe_stash_bytes; tensor path only) for `mlp_jet_bwd_stashed` -- memory instead of recomputation."""
    return _ops.mlp_jet_fwd_stash(x, list(Ws), list(bs), act, jet, precision)


def mlp_jet_bwd_stashed(x: torch.Tensor, Ws: Sequence[torch.Tensor], bs: Sequence[torch.Tensor], act: int, jet: int,
                        u_bar: Optional[torch.Tensor], g_bar: Optional[torch.Tensor], h_bar: Optional[torch.Tensor],
                        stash: torch.Tensor, precision: int = PREC_AUTO):
    """`mlp_jet_bwd` from the stash of `mlp_jet_fwd_stash` (same x, Ws, bs)."""
    out = _ops.mlp_jet_bwd_stashed(x, list(Ws), list(bs), act, jet, u_bar, g_bar, h_bar, stash, precision)
    n = len(Ws)
    return list(out[:n]), list(out[n:])


def mlp_jets(x: torch.Tensor, Ws: Sequence[torch.Tensor], bs: Sequence[torch.Tensor], act: int, jet: int,
             xform=None, precision: int = PREC_AUTO):
    """Differentiable (u [N], g [N,d], h [N,H]).  Gradients w.r.t. the parameters come from the
    reverse kernel; gradients w.r.t. x are assembled from g and h, so nested
    `torch.autograd.grad(u, x, create_graph=True)` calls reproduce the reference idiom.  When the parameters require grad and the stash of all points fits
    (GPE_STASH_MAX_GB, default 128; 80 % of the free device memory), the node keeps it between forward and backward."""
    xa, xb = _xf(xform)
    return _ops.mlp_jets(x, list(Ws), list(bs), act, jet, xa, xb, precision)


class _GramMoments(torch.autograd.Function):
    """m = gram_moments(jets);  backward: (u_bar, g_bar, h_bar, nm_bar) = gram_adjoints(jets, dL/dm)."""

    @staticmethod
    def forward(ctx, u, g, h, nm, x, base, V, q, d, nl_kind, nl_power, scale, bfactor, sym_sign, lap_cols):
        ctx.save_for_backward(u, g, h, nm, x, base, V, q)
        ctx.cfg = (d, nl_kind, nl_power, scale, bfactor, sym_sign, lap_cols)
        return _ops.gram_moments(x, u, g, h, base, V, q, nm, *ctx.cfg)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, c):
        u, g, h, nm, x, base, V, q = ctx.saved_tensors
        ub, gb, hb, nb = _ops.gram_adjoints(x, u, g, h, base, V, q, nm, *ctx.cfg, c.contiguous())
        return (ub, gb if g is not None and g.numel() else None, hb if h is not None and h.numel() else None,
                nb if nm is not None else None) + (None,) * 11


def gram_moments(u, g=None, h=None, nm=None, *, x=None, base=None, V=None, q=None, d=1, nl_kind=NL_NONE, nl_power=3.0,
                 scale=1.0, bfactor=0, sym_sign=0, lap_cols=None):
    """fp64 moment vector m[32] (layout: include/gpe_b200.h) of the full solution
    u = base + scale * F(x) * N built from the network jets (u, g, h); differentiable w.r.t.
    u, g, h and nm."""
    if lap_cols is None:
        lap_cols = 0 if h is None or h.dim() != 2 else (2 if h.shape[1] == 3 else h.shape[1])
    if int(nl_kind) == NL_POW and float(nl_power) != float(int(nl_power)):
        raise ValueError(f"gpe_b200: NL_POW evaluates u**p by repeated multiplication and needs an integer p (got {nl_power}); "
                         "use NL_ABSPOW (|u|^(p-1) u) for a fractional power")
    return _GramMoments.apply(u, g, h, nm, x, base, V, q, int(d), int(nl_kind), float(nl_power), float(scale), int(bfactor),
                              int(sym_sign), int(lap_cols))


def ffma_probe(mode: int, iters: int, device=None):
    """Launch the FFMA peak probe; returns the number of fp32 FLOPs it executes."""
    device = torch.device("cuda") if device is None else device
    sink = torch.zeros(4, dtype=torch.float32, device=device)
    return _ops.ffma_probe(mode, iters, sink)


def set_tensor_path(enable: bool) -> bool:
    """Select the tcgen05 3xTF32 kernels (default where eligible) or the FFMA kernels; returns the previous setting."""
    from . import _lib
    return bool(_lib.lib.gpe_set_tensor_path(1 if enable else 0))


def tc_profile(enable: bool) -> None:
    """Bracket every tensor-path backward kernel with CUDA events (read them with tc_timings())."""
    from . import _lib
    _lib.check(_lib.lib.gpe_tc_profile(1 if enable else 0), "gpe_tc_profile")


def tc_timings():
    """Accumulated milliseconds {stash_forward, reverse_sweep, dw_contraction, fused_backward} since the last call
    (synchronises).  `fused_backward` is the single persistent launch that runs sweep and contraction side by side."""
    import ctypes
    from . import _lib
    ms = (ctypes.c_double * 4)()
    _lib.check(_lib.lib.gpe_tc_timings(ms, 4), "gpe_tc_timings")
    return {"stash_forward": ms[0], "reverse_sweep": ms[1], "dw_contraction": ms[2], "fused_backward": ms[3]}
