"""CUDA-graph replay of a whole training step.

The small configurations of the reference (config 1: 10k points, [1,100,100,100,1]; the paper's curriculum runs:
4000 points, 201 gamma stages x <= 5001 epochs, `R/harmonic_pinn_simulation.py:364-400`) are launch-bound: a step is
~200 small kernels (jet forward, moments, the fp64 scalar algebra of the loss, adjoints, reverse, clip, Adam).  The
ops of this package allocate through torch's caching allocator and launch on the current stream only, so the whole
step -- zero_grad, the reference's loss methods, backward(), clip_grad_norm_, optimizer.step() -- captures into one
CUDA graph.  Measured on B200, config 1 with the root-notebook loop body: 2.2 ms eager -> 0.34-0.41 ms per step (round 1: 0.90).

    step = GraphedStep(lambda: train_body(model, X, ...), optimizer)     # train_body returns tensors (loss, mu, ...)
    for epoch in range(n):
        loss, mu = step()            # replays; the returned tensors are static buffers, overwritten by the next replay

Rules (CUDA graphs): drop every reference to losses / outputs of earlier EAGER steps on the same parameters before
constructing a GraphedStep (a live loss keeps its graph, whose AccumulateGrad nodes stay bound to the eager stream, and
the capture fails with "legacy stream depend on a capturing blocking stream"); tensor shapes and the python control flow of the step must not change between replays; use
`torch.optim.Adam(..., capturable=True)`; call `zero_grad(set_to_none=True)` inside the step (the backward then hands its
gradient tensors over -- with zeroed `.grad` buffers every parameter costs a fill and an accumulate add per step); read results with `.item()`/`float()` only outside the step function
(e.g. every k-th epoch for the early-stop test); scalars that change per stage (gamma of the curriculum) must be
device tensors updated in place with `.fill_()` / `.copy_()`.  Evaluating the model between replays (monitoring every
k-th epoch) is safe: every replay invalidates the models' per-step caches, and evaluations without grad are never cached.
"""
from typing import Callable, Optional

import torch


class GraphedStep:
    """Capture `fn` (one full training step, returning a tensor or a tuple of tensors) once, replay it on call."""

    def __init__(self, fn: Callable, optimizer: Optional[torch.optim.Optimizer] = None, warmup: int = 3):
        if not torch.cuda.is_available():
            raise RuntimeError("GraphedStep needs a CUDA device (gpe_b200 has no CPU path)")
        if optimizer is not None:
            for group in optimizer.param_groups:
                if "capturable" in group and not group["capturable"]:
                    raise ValueError("construct the optimizer with capturable=True to capture its step in a CUDA graph")
        self.fn = fn
        import gc
        gc.collect()      # autograd graphs of earlier eager steps pin their AccumulateGrad nodes to the stream they ran on
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):          # warm-up on a side stream: lazy initialisation, allocator pools, plan caches
            for _ in range(max(warmup, 1)):
                fn()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.outputs = fn()
        self.replays = 0

    def __call__(self):
        self.graph.replay()
        self.replays += 1
        # the replay stepped the parameters on the device without touching Tensor._version: results the models cached
        # for their current autograd graph must not be handed out again (models.invalidate_caches)
        from .models import invalidate_caches
        invalidate_caches()
        return self.outputs
