"""Shared helpers for the parity tests: rebuild oracle models from golden files."""
import numpy as np
import torch


def load_weights(model, g, prefix="w."):
    sd = {k[len(prefix):]: torch.tensor(v) for k, v in g.items() if k.startswith(prefix)}
    model.load_state_dict(sd)
    return model


def golden_grads(g, prefix):
    out, i = [], 0
    while f"{prefix}.{i}" in g:
        out.append(np.asarray(g[f"{prefix}.{i}"]))
        i += 1
    return out


def flat(arrs):
    return np.concatenate([np.asarray(a, dtype=np.float64).ravel() for a in arrs])


def rel_err(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    den = max(np.linalg.norm(b.ravel()), 1e-300)
    return float(np.linalg.norm((a - b).ravel()) / den)


def model_grads(model, total):
    model.zero_grad()
    total.backward()
    return [p.grad.detach().cpu().numpy() if p.grad is not None else np.zeros(tuple(p.shape)) for p in model.parameters()]
