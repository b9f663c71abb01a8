"""Per-point jets (value, gradient, Hessian) of the MLP and the parameter gradient of a linear
jet functional, by nested autograd -- the reference's own derivative idiom
(`torch.autograd.grad(..., create_graph=True)` twice: NBroot cell 6:90-104,
S/gross_pitaevskii_2D.py:183-188, S/helmholtz_2D.py:78-83).

TEST INFRASTRUCTURE (see oracle/__init__.py).  Deliberately *not* Taylor-mode, so it is an
independent check of the CUDA kernels' recurrences.
"""
import torch

ACT_TANH, ACT_SHIFTED_TANH, ACT_LEAKY_RELU, ACT_RESBLOCK_TANH = 0, 1, 2, 3
_EPS = 2.220446049250313e-16


def _act(z, act):
    if act == ACT_TANH:
        return torch.tanh(z)
    if act == ACT_SHIFTED_TANH:
        return torch.tanh(z) + 1.0 + _EPS
    if act == ACT_LEAKY_RELU:
        return torch.nn.functional.leaky_relu(z, 0.01)
    raise ValueError(act)


def mlp(x, Ws, bs, act):
    """x [N,d]; Ws[i] [out,in] (PyTorch layout); returns [N,1]."""
    if act == ACT_RESBLOCK_TANH:
        # R/box_to_gaussian_pinn_simulation.py:52-62,113-130: Linear + ShiftedTanh, then blocks
        # tanh(lin2(tanh(lin1(a))) + a) over consecutive (lin1, lin2) pairs, then the output Linear
        a = _act(x @ Ws[0].t() + bs[0], ACT_SHIFTED_TANH)
        for k in range(1, len(Ws) - 1, 2):
            inner = torch.tanh(a @ Ws[k].t() + bs[k])
            a = torch.tanh(inner @ Ws[k + 1].t() + bs[k + 1] + a)
        return a @ Ws[-1].t() + bs[-1]
    a = x
    for i, (W, b) in enumerate(zip(Ws, bs)):
        a = a @ W.t() + b
        if i < len(Ws) - 1:
            a = _act(a, act)
    return a


def jets(x, Ws, bs, act, dtype=torch.float64):
    """Returns u [N], g [N,d], H [N,d,d] (full Hessian) in `dtype` on CPU."""
    x = x.detach().to("cpu", dtype).clone().requires_grad_(True)
    Ws = [W.detach().to("cpu", dtype) for W in Ws]
    bs = [b.detach().to("cpu", dtype) for b in bs]
    u = mlp(x, Ws, bs, act)
    d = x.shape[1]
    (g,) = torch.autograd.grad(u.sum(), x, create_graph=True)
    rows = []
    for i in range(d):
        if g.requires_grad and g.grad_fn is not None:
            (hi,) = torch.autograd.grad(g[:, i].sum(), x, retain_graph=True, allow_unused=True)
            hi = torch.zeros_like(x) if hi is None else hi
        else:
            hi = torch.zeros_like(x)
        rows.append(hi)
    H = torch.stack(rows, dim=1)
    return u.detach()[:, 0], g.detach(), H.detach()


def jet_functional_param_grads(x, Ws, bs, act, u_bar, g_bar, h_bar, dtype=torch.float64):
    """Gradient w.r.t. every W and b of  sum_n( u_bar*u + <g_bar, g> + <h_bar, H> ).

    h_bar is [N,d,d] (weights on the full Hessian; put the Laplacian weight on the diagonal).
    Returns (dWs, dbs) lists in `dtype`."""
    x = x.detach().to("cpu", dtype).clone().requires_grad_(True)
    Ws = [W.detach().to("cpu", dtype).clone().requires_grad_(True) for W in Ws]
    bs = [b.detach().to("cpu", dtype).clone().requires_grad_(True) for b in bs]
    u = mlp(x, Ws, bs, act)
    d = x.shape[1]
    (g,) = torch.autograd.grad(u.sum(), x, create_graph=True)
    total = (u_bar.to(dtype) * u[:, 0]).sum() + (g_bar.to(dtype) * g).sum()
    for i in range(d):
        if g.grad_fn is None:
            break
        (hi,) = torch.autograd.grad(g[:, i].sum(), x, create_graph=True, allow_unused=True)
        if hi is not None:
            total = total + (h_bar[:, i, :].to(dtype) * hi).sum()
    grads = torch.autograd.grad(total, Ws + bs, allow_unused=True)
    grads = [torch.zeros_like(p) if gr is None else gr for p, gr in zip(Ws + bs, grads)]
    n = len(Ws)
    return grads[:n], grads[n:]
