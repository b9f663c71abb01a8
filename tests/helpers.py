"""Shared helpers for the parity tests: rebuild oracle models from golden files."""
import numpy as np
import torch


def scalar(v):
    """float of a 0-dim / one-element tensor (with or without grad) or array."""
    if torch.is_tensor(v):
        return float(v.detach().reshape(-1)[0])
    return float(np.asarray(v).reshape(-1)[0])


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


VARIANT_GOLDENS = ["vary_harmonic_b0.6_g5_p3_m1", "vary_gaussian_b20_g10_p3_m0", "vary_boxpot_b3_g4_p4_m2",
                   "vary_gravity_shipped_b20_m2", "perturb_shipped_b20_m3", "perturb_b4_gm4_p3_m1"]


def variant_model_and_args(ns, g):
    """(model, pde_loss arguments after (X, u_pred)) of a beta-sweep golden (tests/golden/make_golden.case_variant);
    ns is the module holding the classes: oracle.pinn_oracle or gpe_b200."""
    family, mode, beta, gamma, p = str(g["family"]), int(g["mode"]), float(g["beta"]), float(g["gamma"]), int(g["p"])
    layers, ptype = [int(v) for v in g["layers"]], str(g["ptype"])
    if family == "perturb":
        model = ns.PerturbGammaBetaGPE1D(layers, mode=mode, beta=beta, gamma=gamma)
        args = (p,)
    else:
        model = ns.VaryPotentialGPE1D(layers, mode=mode, beta=beta, family=family)
        args = (gamma, beta, p)
    if ptype:
        args = args + (ptype,)
    return load_weights(model, g), args


OPT_VARIANT_GOLDENS = ["opt_selfadaptive_m0_eta20", "opt_powers_m1_eta30_p4", "opt_powers_m0_eta10_p2", "opt_modes_m2_eta40"]


def opt_variant_model_and_kwargs(ns, g):
    """(model, extra keyword arguments of its loss methods) of an F-opt variant golden (make_golden.case_opt_variant)."""
    kind, layers, mode = str(g["kind"]), [int(v) for v in g["layers"]], int(g["mode"])
    if kind == "self_adaptive":
        model, extra = load_weights(ns.SelfAdaptiveGPE1D(layers, mode=mode), g), {}
        with torch.no_grad():
            model.log_alpha.copy_(torch.tensor(g["log_alpha"]))
    elif kind == "powers":
        model, extra = load_weights(ns.PowersGPE1D(layers, mode=mode), g), dict(p=int(g["p"]))
    else:
        model, extra = load_weights(ns.ModesGPE1D(layers, mode=0), g), dict(mode=mode)
    return model, extra
