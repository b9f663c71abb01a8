"""Data parallelism over collocation points.

The reference is single-process.  Here every rank owns a contiguous block of the collocation set,
the parameters are replicated, and one evaluation of the loss needs exactly two collectives:

  1. SUM all-reduce of the fp64 moment vector (256 B)  -> global mu, norms, energies, means;
  2. SUM all-reduce of the packed parameter gradient (<= 196 KiB) after backward().

Both go through `torch.distributed` (NCCL over NVLink on the GPU box, gloo in the CPU tests).  The
payloads are latency-bound, so they are packed into one buffer each.
"""
import contextlib
import os

import torch
import torch.distributed as dist


def _nvtx(name):
    """Named range around a collective when GPE_NVTX=1 (the kernels' own ranges come from the C library)."""
    if os.environ.get("GPE_NVTX") == "1" and torch.cuda.is_available():
        return torch.cuda.nvtx.range(name)
    return contextlib.nullcontext()


def is_active(group=None):
    return dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1


class _AllReduceSum(torch.autograd.Function):
    """y = sum over ranks of x.  Every rank evaluates the same scalar loss L(y), so
    dL/dx_local = dL/dy: the backward is the identity (no second collective)."""

    @staticmethod
    def forward(ctx, x, group):
        y = x.clone()
        with _nvtx("gpe:all_reduce#1(moments)"):
            dist.all_reduce(y, op=dist.ReduceOp.SUM, group=group)
        return y

    @staticmethod
    def backward(ctx, gy):
        return gy, None


def all_reduce_sum(x: torch.Tensor, group=None) -> torch.Tensor:
    """Differentiable SUM all-reduce for quantities every rank turns into the same global loss
    (moment vectors, quadrature sums).  No-op without an initialised multi-rank group."""
    if not is_active(group):
        return x
    return _AllReduceSum.apply(x, group)


class _ReplicatedTerm(torch.autograd.Function):
    """Identity whose backward divides by the number of ranks."""

    @staticmethod
    def forward(ctx, t, world):
        ctx.world = world
        return t.view_as(t)

    @staticmethod
    def backward(ctx, g):
        return g / ctx.world, None


def replicated_term(t: torch.Tensor, group=None) -> torch.Tensor:
    """Mark a loss term that EVERY rank evaluates on the same (replicated) points -- the boundary / data terms, whose
    2..500 points are not worth sharding.  `all_reduce_gradients` SUMs the ranks' gradients, so a replicated term would
    enter the parameter gradient `world` times while the loss value counts it once; dividing its backward by `world`
    keeps loss and gradient consistent (what an L-BFGS line search needs).  Value unchanged; no-op on one rank."""
    if not is_active(group) or not t.requires_grad:
        return t
    return _ReplicatedTerm.apply(t, float(dist.get_world_size(group)))


def shard_bounds(n: int, rank: int, world: int):
    """Contiguous block [lo, hi) of rank `rank`; block sizes differ by at most one point."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_points(x: torch.Tensor, group=None) -> torch.Tensor:
    """This rank's block of the collocation tensor (rows)."""
    if not is_active(group):
        return x
    lo, hi = shard_bounds(x.shape[0], dist.get_rank(group), dist.get_world_size(group))
    return x[lo:hi]


def all_reduce_gradients(params, group=None):
    """SUM all-reduce of every `.grad` in one packed fp32 buffer (collective #2).  Call after
    backward() and before clip_grad_norm_/optimizer.step(); L-BFGS closures call it before
    returning the loss."""
    if not is_active(group):
        return
    params = [p for p in params if p.grad is not None]
    if not params:
        return
    flat = torch.cat([p.grad.reshape(-1) for p in params])
    with _nvtx("gpe:all_reduce#2(gradient)"):
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    off = 0
    for p in params:
        k = p.grad.numel()
        p.grad.copy_(flat[off:off + k].view_as(p.grad))
        off += k
