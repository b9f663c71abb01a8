"""The paper's curriculum driver on the B200.

`train_gpe_model` mirrors R/harmonic_pinn_simulation.py:220-431 -- same arguments, same return tuple, same loop body
(perturbation scaling, 10 bc + 20 norm + pde, clip_grad_norm_ 1.0, Adam, `CosineAnnealingWarmRestarts.step(total_loss)`,
tolerance / patience early stop, logging every 10 / 100 epochs) -- but a stage (one gamma of one mode: up to 5001
epochs of a 4000-point problem, ~200 small kernels each) is launch-bound, so after the first epochs the whole epoch
is ONE CUDA graph replay and nothing is read back per epoch:

* the loss / mu / constraint of every epoch go into device rings, indexed by a device epoch counter;
* the early-stop tests of :401-412 (`loss <= tol`, 2000 epochs without improvement) run on the device; the parameters
  at the epoch the reference would have left the loop are kept in a snapshot, the host looks at the `done` flag every
  `check_every` epochs and restores that snapshot -- the result is the reference's, whatever `check_every` is;
* the scheduler call of :391 passes the loss where an epoch is expected, which makes the learning rate a function of
  the current loss (torch/optim/lr_scheduler.py, `step(epoch)` branch); that function is evaluated on the device and
  written into Adam's tensor lr (capturable=True).

`best_model_state = model.state_dict().copy()` (:396) is a shallow copy of live tensors, so the reference's final
`load_state_dict(best_model_state)` restores nothing: the stage ends with the last parameters, as here.

`save_models` / `load_models` write and read the reference's pickle bundle (:901-960), so models trained here drop into
the paper's plotting scripts and the shipped `*.pkl` files load into these classes.
"""
import math
import os
import pickle

import numpy as np
import torch
import torch.nn as nn

from .graphs import GraphedStep
from .models import FinalGPE1D

PATIENCE = 2000          # R/harmonic_pinn_simulation.py:361
T_0, T_MULT, ETA_MIN = 200, 2, 1e-6   # :347-349


def advanced_initialization(m, mode):
    """:636-648."""
    if isinstance(m, nn.Linear):
        nn.init.xavier_normal_(m.weight, gain=1.0 / (1.0 + 0.2 * mode))
        m.bias.data.fill_(0.001 if mode > 3 else 0.01)


def pretrain_on_analytical_solution(model, mode, X_train, epochs=5000, lr=1e-3, verbose=False):
    """:650-702: fit the network to psi_n (Adam, the last 500 epochs L-BFGS).  Value-only jets; the per-epoch
    `loss.item()` of the reference is kept (it is the convergence test)."""
    model.train()
    dev = next(model.parameters()).device
    X = torch.tensor(X_train, dtype=torch.float32, device=dev)
    with torch.no_grad():
        target = model.weighted_hermite(X, mode).detach()
    opt, closure, loss = torch.optim.Adam(model.parameters(), lr=lr), None, None
    for epoch in range(epochs):
        if epoch < epochs - 500:
            opt.zero_grad()
            loss = torch.mean((model.forward(X) - target) ** 2)
            loss.backward()
            opt.step()
        else:
            if epoch == epochs - 500:
                opt = torch.optim.LBFGS(model.parameters(), lr=lr * 0.1, max_iter=20)

                def closure():
                    opt.zero_grad()
                    val = torch.mean((model.forward(X) - target) ** 2)
                    val.backward()
                    return val
            loss = opt.step(closure)
        if verbose and epoch % 500 == 0:
            print(f"  Pre-training epoch {epoch}, loss: {loss.item():.2e}")
        if loss.item() < 1e-12:
            break
    return model


def _loss_as_epoch_lr(loss, base_lr):
    """CosineAnnealingWarmRestarts(T_0=200, T_mult=2, eta_min=1e-6).step(loss) as tensor arithmetic."""
    n = torch.floor(torch.log2(torch.clamp(loss / T_0 * (T_MULT - 1) + 1.0, min=1.0)))
    restart = loss >= T_0
    t_cur = torch.where(restart, loss - T_0 * (torch.pow(2.0, n) - 1.0) / (T_MULT - 1), loss)
    t_i = torch.where(restart, T_0 * torch.pow(2.0, n), torch.full_like(loss, float(T_0)))
    return ETA_MIN + (base_lr - ETA_MIN) * (1.0 + torch.cos(math.pi * t_cur / t_i)) / 2.0


class _Stage:
    """One (mode, gamma) stage: device-side bookkeeping + the epoch body."""

    def __init__(self, model, X, bp, bv, dx, gamma, p, potential_type, perturb_const, lr, epochs, tol, pde_args=None):
        dev = X.device
        self.model, self.X, self.bp, self.bv, self.dx = model, X, bp, bv, dx
        self.gamma, self.p, self.ptype, self.q, self.tol = gamma, p, potential_type, perturb_const, tol
        self.pde_args = pde_args
        self.base_lr = lr
        self.lr = torch.tensor(float(lr), device=dev)                  # scheduler construction leaves lr = base_lr
        self.opt = torch.optim.Adam(model.parameters(), lr=self.lr, capturable=True)
        self.params = [q for q in model.parameters()]
        self.snap = [torch.zeros_like(q) for q in self.params]
        self.epoch = torch.zeros((), dtype=torch.long, device=dev)
        self.final_epoch = torch.full((), epochs, dtype=torch.long, device=dev)
        self.done = torch.zeros((), dtype=torch.bool, device=dev)
        self.best = torch.full((), float("inf"), device=dev)
        self.stale = torch.zeros((), dtype=torch.long, device=dev)
        self.rings = torch.zeros(3, epochs, device=dev)               # total, mu, constraint per epoch
        self.normal_const = None

    def body(self, first_of_mode=False):
        m = self.model
        self.opt.zero_grad(set_to_none=True)     # the backward hands its gradient tensors over: no fill, no accumulate add
        u_pred = m.forward(self.X)
        if first_of_mode:                                              # :366-371
            self.normal_const = float(torch.max(u_pred).detach())
            u_pred = self.q * (u_pred / self.normal_const)
        else:
            u_pred = (self.q * u_pred) / self.normal_const
        bc = m.boundary_loss(self.bp, self.bv)
        nrm = m.normalization_loss(m.get_complete_solution(self.X, u_pred), self.dx)
        cons = 10.0 * bc + 20.0 * nrm
        args = self.pde_args if self.pde_args is not None else (self.gamma, self.p, self.ptype)
        pde, mu = m.pde_loss(self.X, u_pred, *args)
        total = pde + cons
        total.backward()
        torch.nn.utils.clip_grad_norm_(self.params, 1.0)
        self.opt.step()
        with torch.no_grad():
            # scheduler.step(loss) (:391), logs, patience (:394-399), early stop and snapshot (:401-412): one launch on device scalars
            val = total.detach().float().reshape(())
            torch.ops.gpe.stage_bookkeeping(val, mu.detach().float().reshape(()), cons.detach().float().reshape(()), self.lr, self.epoch,
                                            self.final_epoch, self.done, self.best, self.stale, self.rings, self.params, self.snap,
                                            float(self.base_lr), float(self.tol), PATIENCE, float(T_0), float(T_MULT), float(ETA_MIN))
        return val

    def finish(self, ran):
        """Host side, once per stage: restore the parameters of the stop epoch, cut the logs as :414-423 does."""
        stopped = bool(self.done)
        last = int(self.final_epoch) if stopped else ran - 1
        if stopped:
            with torch.no_grad():
                for s, q in zip(self.snap, self.params):
                    q.copy_(s)
        rings = self.rings[:, : last + 1].cpu().numpy()
        hist = {"loss": [float(v) for v in rings[0, ::10]], "constraint": [float(v) for v in rings[2, ::100]],
                "lambda": [float(v) for v in rings[1, ::100]]}
        return hist, (last if stopped else ran), rings


def train_gpe_model(gamma_values, modes, p, X_train, lb, ub, layers, epochs, tol, perturb_const, potential_type="harmonic",
                    lr=1e-5, verbose=True, *, model_factory=None, pretrain_epochs=2000, check_every=50, use_graph=True,
                    init_state=None, device="cuda", keep_rings=False, normal_const=None):
    """R/harmonic_pinn_simulation.py:220-431 -> (models_by_mode, mu_table, training_history, constant_history,
    epochs_history).  Keyword-only extras: model_factory(layers, mode, gamma) (default FinalGPE1D of `potential_type`),
    pretrain_epochs (:338 uses 2000), check_every (how often the host reads the device `done` flag), use_graph,
    init_state (state_dict to start every mode from instead of pre-training / advanced_initialization),
    keep_rings (also return the per-epoch device logs in training_history[mode][gamma]['per_epoch']),
    normal_const (float or {mode: float}: continue a run whose normalisation constant is known -- the `constant_history`
    of a shipped bundle -- instead of fixing it at epoch 0 of the first stage)."""
    if not torch.cuda.is_available():
        raise RuntimeError("gpe_b200.training needs a CUDA device (there is no CPU path)")
    dx = float(X_train[1, 0] - X_train[0, 0])
    X = torch.tensor(X_train, dtype=torch.float32, device=device).requires_grad_(True)
    bp = torch.tensor([[lb], [ub]], dtype=torch.float32, device=device)
    bv = torch.zeros((2, 1), dtype=torch.float32, device=device)
    if model_factory is None:
        def model_factory(layers_, mode_, gamma_):
            return FinalGPE1D(layers_, mode=mode_, gamma=gamma_, potential=potential_type)
    models_by_mode, mu_table, training_history, constant_history, epochs_history = {}, {}, {}, {}, {}
    gamma_values = sorted(gamma_values)
    for mode in modes:
        mu_logs, by_gamma, hist_by_gamma, epochs_by_gamma, prev = [], {}, {}, {}, None
        nc_given = normal_const.get(mode) if isinstance(normal_const, dict) else normal_const
        normal_const_m = float(nc_given) if nc_given is not None else None
        if normal_const_m is not None:
            constant_history[mode] = torch.tensor(normal_const_m)
        for gamma in gamma_values:
            model = model_factory(layers, mode, gamma).to(device)
            if prev is not None:
                model.load_state_dict(prev.state_dict())
            elif init_state is not None:
                model.load_state_dict(init_state)
            elif gamma == 0.0:
                model = pretrain_on_analytical_solution(model, mode, X_train, epochs=pretrain_epochs, lr=1e-3, verbose=verbose)
            else:
                model.apply(lambda m: advanced_initialization(m, mode))
            st = _Stage(model, X, bp, bv, dx, gamma, p, potential_type, perturb_const, lr, epochs, tol)
            st.normal_const = normal_const_m
            # epoch 0 of a mode's first stage fixes the normalisation constant (:366-371); it runs eagerly
            first = normal_const_m is None
            if first and gamma != 0:
                raise ValueError("the first gamma of a mode must be 0: the reference sets normal_const there (:366)")
            st.body(first_of_mode=first)
            ran = 1
            normal_const_m = st.normal_const
            if first:
                constant_history[mode] = torch.tensor(normal_const_m)
            step = st.body
            if use_graph and epochs - ran > 8:
                step = GraphedStep(st.body, st.opt, warmup=2)          # two more eager epochs, then one capture
                ran += 2
            while ran < epochs:
                if ran % check_every == 0 and bool(st.done):           # the only per-stage host reads
                    break
                step()
                ran += 1
            hist, final_epoch, rings = st.finish(ran)
            if keep_rings:
                hist["per_epoch"] = rings
            final_mu = hist["lambda"][-1] if hist["lambda"] else 0
            mu_logs.append((gamma, final_mu))
            by_gamma[gamma], hist_by_gamma[gamma], epochs_by_gamma[gamma] = model, hist, final_epoch
            prev = model
            if verbose:
                print(f"mode {mode} gamma {gamma}: epochs {final_epoch}, mu {final_mu:.6f}, loss {hist['loss'][-1]:.3e}")
        models_by_mode[mode], mu_table[mode] = by_gamma, mu_logs
        training_history[mode], epochs_history[mode] = hist_by_gamma, epochs_by_gamma
    return models_by_mode, mu_table, training_history, constant_history, epochs_history


def save_models(models_by_mode, mu_table, training_history, constant_history, epochs_history, filename="gpe_models.pkl",
                save_dir="."):
    """:901-929: the reference's bundle (CPU state_dicts + constructor fields); models stay on their device."""
    bundle = {}
    for mode, by_gamma in models_by_mode.items():
        bundle[mode] = {}
        for gamma, model in by_gamma.items():
            bundle[mode][gamma] = {"state_dict": {k: v.detach().cpu() for k, v in model.state_dict().items()},
                                   "layers": model.layers, "hbar": model.hbar, "m": model.m, "mode": model.mode,
                                   "gamma": model.gamma}
    data = {"models_state_dicts": bundle, "mu_table": mu_table,
            "training_history": {m: {g: {k: v for k, v in h.items() if k != "per_epoch"} for g, h in hg.items()}
                                 for m, hg in training_history.items()},
            "constant_history": constant_history, "epochs_history": epochs_history}
    os.makedirs(save_dir, exist_ok=True)
    path = os.path.join(save_dir, filename)
    with open(path, "wb") as fh:
        pickle.dump(data, fh)
    return path


def load_models(filename="gpe_models.pkl", save_dir=".", *, model_factory=None, potential="harmonic", device="cuda"):
    """:931-960: rebuild the models of a bundle (also the reference's shipped `*_plot_data.pkl` files)."""
    with open(os.path.join(save_dir, filename), "rb") as fh:
        data = pickle.load(fh)
    models_by_mode = {}
    for mode, by_gamma in data["models_state_dicts"].items():
        models_by_mode[mode] = {}
        for gamma, md in by_gamma.items():
            if model_factory is not None:
                model = model_factory(md["layers"], md["mode"], md["gamma"])
            else:
                model = FinalGPE1D(md["layers"], hbar=md["hbar"], m=md["m"], mode=md["mode"], gamma=md["gamma"],
                                   potential=potential)
            model.load_state_dict(md["state_dict"])
            models_by_mode[mode][gamma] = model.to(device).eval()
    return (models_by_mode, data["mu_table"], data["training_history"], data["constant_history"], data["epochs_history"])
