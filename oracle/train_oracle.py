"""TEST INFRASTRUCTURE (see oracle/__init__.py): CPU restatement of the paper's curriculum driver
`train_gpe_model` (R/harmonic_pinn_simulation.py:220-431) for the oracle classes, any dtype.  Uses torch's own
`CosineAnnealingWarmRestarts` exactly as the reference does (`scheduler.step(total_loss)`, :391), so the B200 driver's
device-side restatement of that call is checked against the real scheduler.  Pre-training (:338) is left out: tests start
both drivers from the same `init_state`."""
import torch
from torch.optim.lr_scheduler import CosineAnnealingWarmRestarts

from . import pinn_oracle as po


def train_gpe_model(gamma_values, modes, p, X_train, lb, ub, layers, epochs, tol, perturb_const, potential_type="harmonic",
                    lr=1e-5, init_state=None, dtype=torch.float64, patience=2000):
    dx = float(X_train[1, 0] - X_train[0, 0])
    X = torch.tensor(X_train, dtype=dtype, requires_grad=True)
    bp = torch.tensor([[lb], [ub]], dtype=dtype)
    bv = torch.zeros((2, 1), dtype=dtype)
    models_by_mode, mu_table, training_history, constant_history, epochs_history = {}, {}, {}, {}, {}
    for mode in modes:
        mu_logs, by_gamma, hist_by_gamma, epochs_by_gamma, prev, normal_const = [], {}, {}, {}, None, None
        for gamma in sorted(gamma_values):
            model = po.FinalGPE1D(layers, mode=mode, gamma=gamma, potential=potential_type).to(dtype)
            model.load_state_dict(prev.state_dict() if prev is not None else init_state)
            opt = torch.optim.Adam(model.parameters(), lr=lr)
            sched = CosineAnnealingWarmRestarts(opt, T_0=200, T_mult=2, eta_min=1e-6)
            lam_h, loss_h, cons_h, per_epoch = [], [], [], []
            best, stale, final_epoch = float("inf"), 0, epochs
            for epoch in range(epochs):
                opt.zero_grad()
                u_pred = model.forward(X)
                if epoch == 0 and gamma == 0:
                    normal_const = torch.max(u_pred).detach().clone()
                    constant_history[mode] = normal_const
                    u_pred = perturb_const * (u_pred / normal_const)
                else:
                    u_pred = (perturb_const * u_pred) / normal_const
                bc = model.boundary_loss(bp, bv)
                nrm = model.normalization_loss(model.get_complete_solution(X, u_pred), dx)
                cons = 10.0 * bc + 20.0 * nrm
                pde, mu = model.pde_loss(X, u_pred, gamma, p, potential_type)
                total = pde + cons
                total.backward()
                torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
                opt.step()
                sched.step(total)
                cur = total.item()
                per_epoch.append((cur, mu.item(), cons.item(), opt.param_groups[0]["lr"]))
                if cur < best:
                    best, stale = cur, 0
                else:
                    stale += 1
                if epoch % 10 == 0:
                    loss_h.append(cur)
                if epoch % 100 == 0:
                    lam_h.append(mu.item())
                    cons_h.append(cons.item())
                if cur <= tol or stale >= patience:
                    final_epoch = epoch
                    break
            mu_logs.append((gamma, lam_h[-1] if lam_h else 0))
            by_gamma[gamma] = model
            hist_by_gamma[gamma] = {"loss": loss_h, "constraint": cons_h, "lambda": lam_h, "per_epoch": per_epoch}
            epochs_by_gamma[gamma] = final_epoch
            prev = model
        models_by_mode[mode], mu_table[mode] = by_gamma, mu_logs
        training_history[mode], epochs_history[mode] = hist_by_gamma, epochs_by_gamma
    return models_by_mode, mu_table, training_history, constant_history, epochs_history
