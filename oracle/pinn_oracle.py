"""Plain-PyTorch CPU restatement of the reference PINN classes (test oracle / CPU baseline).

TEST INFRASTRUCTURE: see oracle/__init__.py.  Every class keeps the reference's method names,
argument order, return arity, state_dict keys and tensor shapes (including the reference's
`[N]` x `[N,1]` broadcast in the 2-D files), but is written independently and runs in whatever
dtype its parameters have, so the same code serves as the fp32 "port" baseline and the fp64
truth.

Short path names (all under /root/reference/):
  NBroot = Gross_Pitaevskii_1D_power_Test.ipynb          (cell N:line L)
  R/     = Gross-Pitaevskii/src/final/refine/
  S/     = Gross-Pitaevskii/src/
"""
import math

import numpy as np
import torch
import torch.nn as nn
from torch.autograd import grad as _grad


# --------------------------------------------------------------------------------------
# activations and the MLP
# --------------------------------------------------------------------------------------
class ShiftedTanh(nn.Module):
    """tanh(x) + 1 + eps  (R/harmonic_pinn_simulation.py:41-49)."""

    def __init__(self, eps=float(np.finfo(float).eps)):
        super().__init__()
        self.eps = eps

    def forward(self, z):
        return torch.tanh(z) + 1.0 + self.eps


def sequential_mlp(widths, act_factory):
    """Linear/act/.../Linear as nn.Sequential; keys network.{0,2,4,...}
    (NBroot cell 6:30-39, R/harmonic_pinn_simulation.py:84-93, S/gross_pitaevskii_2D.py:59-65)."""
    mods = []
    last = len(widths) - 2
    for i in range(len(widths) - 1):
        mods.append(nn.Linear(widths[i], widths[i + 1]))
        if i < last:
            mods.append(act_factory())
    return nn.Sequential(*mods)


def d_dx(out, x):
    """`torch.autograd.grad(out, x, ones_like(out), create_graph=True)[0]`, the reference's
    derivative idiom (NBroot cell 6:90-104)."""
    return _grad(out, x, grad_outputs=torch.ones_like(out), create_graph=True, retain_graph=True)[0]


def xavier_bias001_(module):
    """S/gross_pitaevskii_2D.py:614-618 / S/gross_pitaevskii_1D_Different_Optimizers.py:443-455."""
    if isinstance(module, nn.Linear):
        nn.init.xavier_uniform_(module.weight)
        module.bias.data.fill_(0.01)


# --------------------------------------------------------------------------------------
# analytic base states
# --------------------------------------------------------------------------------------
def hermite_poly_np(n, x):
    """Physicists' Hermite polynomial H_n by the three-term recurrence (numpy, float64)."""
    x = np.asarray(x, dtype=np.float64)
    h0 = np.ones_like(x)
    if n == 0:
        return h0
    h1 = 2.0 * x
    for k in range(1, n):
        h0, h1 = h1, 2.0 * x * h1 - 2.0 * k * h0
    return h1


def hermite_norm(n):
    return (2.0 ** n * math.factorial(n) * math.sqrt(math.pi)) ** (-0.5)


def hermite_state_constH(x, n):
    """NBroot cell 6:41-48 / S/gross_pitaevskii_1D_Different_Optimizers.py:89-111: H_n is
    evaluated on a *detached* numpy copy of x (so it is a constant for autograd) and cast to
    float32; only exp(-x^2/2) is differentiated."""
    h = hermite_poly_np(n, x.detach().cpu().numpy())
    h_t = torch.tensor(h, dtype=torch.float32).to(x.device)
    return hermite_norm(n) * torch.exp(-x ** 2 / 2) * h_t


def hermite_state_torch(x, n):
    """R/harmonic_pinn_simulation.py:95-119: recurrence in torch, fully differentiable."""
    if n == 0:
        hn = torch.ones_like(x)
    elif n == 1:
        hn = 2.0 * x
    else:
        a, b = torch.ones_like(x), 2.0 * x
        for k in range(1, n):
            hn = 2.0 * x * b - 2.0 * float(k) * a
            a, b = b, hn
    return torch.tensor(hermite_norm(n), dtype=x.dtype, device=x.device) * (hn * torch.exp(-0.5 * x ** 2))


# --------------------------------------------------------------------------------------
# F-root: the root notebook class (BASELINE config 1)
# --------------------------------------------------------------------------------------
class RootGPE1D(nn.Module):
    """NBroot cell 6 `GrossPitaevskiiPINN(layers, hbar, m, mode, gamma, power)`."""

    def __init__(self, layers, hbar=1.0, m=1.0, mode=0, gamma=1.0, power=1.0):
        super().__init__()
        self.layers = layers
        self.network = sequential_mlp(layers, nn.Tanh)
        self.hbar, self.m, self.mode, self.gamma, self.power = hbar, m, mode, gamma, power

    def weighted_hermite(self, x, n):
        return hermite_state_constH(x, n)

    def forward(self, inputs):
        return self.network(inputs)

    def get_complete_solution(self, x, perturbation, mode=0):
        return self.weighted_hermite(x, mode) + perturbation

    def compute_potential(self, x, potential_type="harmonic", **kw):
        """NBroot cell 6:63-79."""
        if potential_type == "harmonic":
            return 0.5 * kw.get("omega", 1.0) ** 2 * x ** 2
        if potential_type == "gaussian":
            return torch.exp(-(x - kw.get("a", 0.0)) ** 2)
        if potential_type == "periodic":
            return kw.get("V0", 1.0) * torch.cos(kw.get("k", 2 * np.pi / 5.0) * x) ** 2
        raise ValueError(f"Unknown potential type: {potential_type}")

    def pde_loss(self, p, inputs, predictions, gamma, potential_type="harmonic", precomputed_potential=None):
        """NBroot cell 6:81-127: Hu = -u''/2 + V u + gamma u**p, Rayleigh mu, mean r^2."""
        u = self.get_complete_solution(inputs, predictions, self.mode)
        u_x = d_dx(u, inputs)
        u_xx = d_dx(u_x, inputs)
        V = precomputed_potential if precomputed_potential is not None else self.compute_potential(inputs, potential_type)
        hu = -0.5 * u_xx + V * u + gamma * u ** p
        mu = torch.mean(u * hu) / torch.mean(u ** 2)
        r = hu - mu * u
        return torch.mean(r ** 2), r, mu, u

    def boundary_loss(self, boundary_points, boundary_values):
        """NBroot cell 6:129-135."""
        full = self.get_complete_solution(boundary_points, self.forward(boundary_points), self.mode)
        return torch.mean((full - boundary_values) ** 2)

    def symmetry_loss(self, collocation_points, lb, ub):
        """NBroot cell 6:137-155: NN output only, mirror about 0, parity from the mode."""
        a = self.forward(collocation_points)
        b = self.forward(-collocation_points)
        return torch.mean((a + b) ** 2) if self.mode % 2 == 1 else torch.mean((a - b) ** 2)

    def normalization_loss(self, u, dx):
        """NBroot cell 6:157-162."""
        return (torch.sum(u ** 2) * dx - 1.0) ** 2


def root_train_terms(model, X, boundary_points, boundary_values, dx, gamma, power, lb, ub,
                     potential_type="harmonic"):
    """Loop body NBroot cell 10:88-97 (without the optimizer): returns
    (total, pde, bc, norm, sym, mu)."""
    u_pred = model.forward(X)
    pde, _, mu, full_u = model.pde_loss(power, X, u_pred, gamma, potential_type)
    bc = model.boundary_loss(boundary_points, boundary_values)
    nrm = model.normalization_loss(full_u, dx)
    sym = model.symmetry_loss(X, lb, ub)
    return pde + 10.0 * bc + 20.0 * nrm + 5.0 * sym, pde, bc, nrm, sym, mu


class PaperGPE1D(RootGPE1D):
    """NBP/Gross_Pitaevskii_1D_Harmonic.ipynb cell 6 `GrossPitaevskiiPINN(layers, hbar, m, mode, gamma)`: the root
    notebook's network and constant-H Hermite base with gamma |u|^(p-1) u, mu returned as mu + mode, a normalised
    Riesz energy and the symmetry term on the full solution."""

    def __init__(self, layers, hbar=1.0, m=1.0, mode=0, gamma=1.0):
        super().__init__(layers, hbar, m, mode, gamma)

    def get_complete_solution(self, x, perturbation, mode=None):
        return self.weighted_hermite(x, self.mode if mode is None else mode) + perturbation

    def pde_loss(self, inputs, predictions, gamma, p, potential_type="harmonic", precomputed_potential=None):
        """cell 6:82-131 -> (pde_loss, pde_residual, lambda_pde + mode, u)."""
        u = self.get_complete_solution(inputs, predictions)
        u_xx = d_dx(d_dx(u, inputs), inputs)
        V = precomputed_potential if precomputed_potential is not None else self.compute_potential(inputs, potential_type)
        hu = -0.5 * u_xx + V * u + gamma * torch.abs(u) ** (p - 1) * u
        mu = torch.mean(u * hu) / torch.mean(u ** 2)
        r = hu - mu * u
        return torch.mean(r ** 2), r, mu + self.mode, u

    def riesz_loss(self, inputs, predictions, gamma, p, potential_type="harmonic", precomputed_potential=None):
        """cell 6:133-183 -> (E, E + mode, u), E = (1/2 S u'^2 + S V u^2 + gamma/(p+1) S |u|^(p+1)) / S u^2 with
        S = sum * dx, dx = inputs[1] - inputs[0]."""
        u = self.get_complete_solution(inputs, predictions)
        if not inputs.requires_grad:
            inputs = inputs.clone().detach().requires_grad_(True)
        u_x = d_dx(u, inputs)
        dx = inputs[1] - inputs[0]
        nf = torch.sum(u ** 2) * dx
        V = precomputed_potential if precomputed_potential is not None else self.compute_potential(inputs, potential_type)
        e = (0.5 * torch.sum(u_x ** 2) * dx / nf + torch.sum(V * u ** 2) * dx / nf
             + 0.5 * (2.0 * gamma / (p + 1)) * torch.sum(torch.abs(u) ** (p + 1)) * dx / nf)
        return e, e + self.mode, u

    def boundary_loss(self, boundary_points, boundary_values):
        full = self.get_complete_solution(boundary_points, self.forward(boundary_points))
        return torch.mean((full - boundary_values) ** 2)

    def symmetry_loss(self, collocation_points, lb, ub):
        """cell 6:193-214: on base + network."""
        xr = -collocation_points
        a = self.get_complete_solution(collocation_points, self.forward(collocation_points))
        b = self.get_complete_solution(xr, self.forward(xr))
        return torch.mean((a + b) ** 2) if self.mode % 2 == 1 else torch.mean((a - b) ** 2)


def paper_train_terms(model, X, boundary_points, boundary_values, dx, gamma, p, lb, ub, potential_type="harmonic"):
    """Loop body NBP/Gross_Pitaevskii_1D_Harmonic.ipynb cell 8:97-145: mode 0 minimises Riesz + pde, the others pde;
    returns (total, physics, bc, norm, sym, lambda_value)."""
    u_pred = model.forward(X)
    bc = model.boundary_loss(boundary_points, boundary_values)
    nrm = model.normalization_loss(model.get_complete_solution(X, u_pred), dx)
    sym = model.symmetry_loss(X, lb, ub)
    if model.mode == 0:
        e, lam, _ = model.riesz_loss(X, u_pred, gamma, p, potential_type)
        pde, _, _, _ = model.pde_loss(X, u_pred, gamma, p, potential_type)
        phys = e + pde
    else:
        phys, _, lam, _ = model.pde_loss(X, u_pred, gamma, p, potential_type)
    return phys + 10.0 * bc + 20.0 * nrm + 5.0 * sym, phys, bc, nrm, sym, lam


# --------------------------------------------------------------------------------------
# F-final: the paper's PL-PINN scripts (own the shipped .pkl fixtures)
# --------------------------------------------------------------------------------------
class FinalGPE1D(nn.Module):
    """R/harmonic_pinn_simulation.py:52-217 `GrossPitaevskiiPINN(layers, hbar, m, mode, gamma,
    use_perturbation)`; `potential='box'` follows R/box_pinn_simulation.py:52-262 (sin(pi x)
    boundary factor in forward, analytic base derivatives); `potential='gravity_well'` follows
    R/gravity_well_pinn_simulation.py:53-256 (Airy base from scipy on the host, normalised on the grid it is
    given, its second derivative by np.gradient; V = x)."""

    def __init__(self, layers, hbar=1.0, m=1.0, mode=0, gamma=1.0, use_perturbation=True,
                 potential="harmonic", L=1.0, enforce_boundary=None):
        super().__init__()
        self.layers = layers
        self.network = sequential_mlp(layers, ShiftedTanh)
        self.hbar, self.m, self.mode, self.gamma = hbar, m, mode, gamma
        self.use_perturbation = use_perturbation
        self.potential, self.L = potential, L
        # R/plot_box_potential_at_ground_state.py:58-91,124-137 makes the sin(pi x) factor of the box forward an
        # option (False for its vanilla PINN); None = the box script's behaviour (always on)
        self.enforce_boundary = (potential == "box") if enforce_boundary is None else enforce_boundary

    def weighted_hermite(self, x, n):
        return hermite_state_torch(x, n)

    def box_eigenfunction(self, x, n):
        """R/box_pinn_simulation.py:99-117."""
        return torch.sqrt(torch.tensor(2.0 / self.L)) * torch.sin((n + 1) * torch.pi * x / self.L)

    def airy_solution(self, x, n):
        """R/gravity_well_pinn_simulation.py:97-118: Ai(x + alpha_n) / sqrt(sum(Ai^2) dx), dx from the points given
        (also for the two boundary points), float32 result whatever the dtype of x."""
        from scipy.special import ai_zeros, airy
        x_np = x.detach().cpu().numpy()
        alpha_n = ai_zeros(n + 1)[0][n]
        psi = airy(x_np + alpha_n)[0]
        dx = x_np[1] - x_np[0] if len(x_np) > 1 else 0.01
        psi = psi / np.sqrt(np.sum(psi ** 2) * dx)
        return torch.tensor(psi, dtype=torch.float32).to(x.device)

    def forward(self, inputs):
        out = self.network(inputs)
        if self.enforce_boundary:                        # R/box_pinn_simulation.py:119-130
            out = out * torch.sin(torch.pi * inputs)
        return out

    def get_complete_solution(self, x, perturbation, mode=None):
        n = self.mode if mode is None else mode
        if self.potential == "gravity_well":
            return self.airy_solution(x, n) + perturbation
        base = self.box_eigenfunction(x, n) if self.potential == "box" else self.weighted_hermite(x, n)
        return base + perturbation

    def compute_potential(self, x, potential_type="harmonic", **kw):
        if potential_type == "harmonic":                 # R/harmonic_pinn_simulation.py:136-144
            return x ** 2
        if potential_type == "box":                      # R/box_pinn_simulation.py:182-190
            return torch.zeros_like(x)
        if potential_type == "gravity_well":             # R/gravity_well_pinn_simulation.py:175-184
            return x.clone()
        raise ValueError(f"Unknown potential type: {potential_type}")

    def _jets(self, inputs, predictions):
        if not self.use_perturbation:
            u = predictions
            u_x = d_dx(u, inputs)
            return u, u_x, d_dx(u_x, inputs)
        if self.potential == "box":                      # R/box_pinn_simulation.py:141-180
            k = (self.mode + 1) * np.pi / self.L
            xn = inputs.detach().cpu().numpy().flatten()
            c = np.sqrt(2.0 / self.L)
            as_col = lambda a: torch.tensor(a, dtype=torch.float32).reshape(-1, 1).to(inputs.device)
            b0, b1, b2 = as_col(c * np.sin(k * xn)), as_col(c * k * np.cos(k * xn)), as_col(-c * k ** 2 * np.sin(k * xn))
            p_x = d_dx(predictions, inputs)
            p_xx = d_dx(p_x, inputs)
            return b0 + predictions, b1 + p_x, b2 + p_xx
        if self.potential == "gravity_well":             # R/gravity_well_pinn_simulation.py:135-173
            from scipy.special import ai_zeros, airy
            xn = inputs.detach().cpu().numpy().flatten()
            ai_val, aip_val, _, _ = airy(xn + ai_zeros(self.mode + 1)[0][self.mode])
            dx = xn[1] - xn[0] if len(xn) > 1 else 0.01
            nf = np.sqrt(np.sum(ai_val ** 2) * dx)
            as_col = lambda a: torch.tensor(a, dtype=torch.float32).reshape(-1, 1).to(inputs.device)
            b0, b1, b2 = as_col(ai_val / nf), as_col(aip_val / nf), as_col(np.gradient(aip_val / nf, dx))
            p_x = d_dx(predictions, inputs)
            p_xx = d_dx(p_x, inputs)
            return b0 + predictions, b1 + p_x, b2 + p_xx
        u = self.get_complete_solution(inputs, predictions)
        u_x = d_dx(u, inputs)
        return u, u_x, d_dx(u_x, inputs)

    def pde_loss(self, inputs, predictions, gamma, p, potential_type=None, precomputed_potential=None):
        """R/harmonic_pinn_simulation.py:146-196: Hu = -u'' + V u + gamma u**p.  (The gravity-well script forms
        mu as sum/sum, :227-229 -- the same number.)"""
        potential_type = potential_type or self.potential
        u, _, u_xx = self._jets(inputs, predictions)
        V = precomputed_potential if precomputed_potential is not None else self.compute_potential(inputs, potential_type)
        hu = -u_xx + V * u + gamma * u ** p
        mu = torch.mean(u * hu) / torch.mean(u ** 2)
        return torch.mean((hu - mu * u) ** 2), mu

    def boundary_loss(self, boundary_points, boundary_values):
        """R/harmonic_pinn_simulation.py:198-210 (unscaled NN output + base)."""
        out = self.forward(boundary_points)
        full = self.get_complete_solution(boundary_points, out) if self.use_perturbation else out
        return torch.mean((full - boundary_values) ** 2)

    def normalization_loss(self, u, dx):
        return (torch.sum(u ** 2) * dx - 1.0) ** 2


def final_train_terms(model, X, boundary_points, boundary_values, dx, gamma, p, perturb_const, normal_const):
    """Loop body R/harmonic_pinn_simulation.py:332-355: returns (total, pde, bc, norm, mu)."""
    u_pred = perturb_const * model.forward(X) / normal_const
    bc = model.boundary_loss(boundary_points, boundary_values)
    nrm = model.normalization_loss(model.get_complete_solution(X, u_pred), dx)
    pde, mu = model.pde_loss(X, u_pred, gamma, p)
    return pde + 10.0 * bc + 20.0 * nrm, pde, bc, nrm, mu


class ResidualBlock(nn.Module):
    """R/box_to_gaussian_pinn_simulation.py:52-62."""

    def __init__(self, dim):
        super().__init__()
        self.lin1 = nn.Linear(dim, dim)
        self.lin2 = nn.Linear(dim, dim)

    def forward(self, x):
        return torch.tanh(self.lin2(torch.tanh(self.lin1(x))) + x)


def residual_mlp(widths):
    """R/box_to_gaussian_pinn_simulation.py:113-130: Linear + ShiftedTanh, len(widths) - 3 residual blocks, Linear."""
    mods = [nn.Linear(widths[0], widths[1]), ShiftedTanh()]
    mods += [ResidualBlock(widths[1]) for _ in range(len(widths) - 3)]
    mods.append(nn.Linear(widths[1], widths[-1]))
    return nn.Sequential(*mods)


class BoxToGaussianGPE1D(FinalGPE1D):
    """R/box_to_gaussian_pinn_simulation.py:65-247 `GrossPitaevskiiPINN(layers, hbar, m, mode, gamma, L, use_residual)`:
    residual-block network, box-sine base differentiated by autograd (no analytic bundle, no sin(pi x) factor),
    V = exp(-(x - 1/2)^2), Hu = -u'' + V u + gamma u**p."""

    def __init__(self, layers, hbar=1.0, m=1.0, mode=0, gamma=1.0, L=1.0, use_residual=True):
        super().__init__(layers, hbar, m, mode, gamma, True, potential="box", L=L, enforce_boundary=False)
        self.use_residual = use_residual
        if use_residual:
            self.network = residual_mlp(layers)

    def compute_potential(self, x, potential_type="gaussian", **kw):
        if potential_type != "gaussian":                 # :167-175
            raise ValueError(f"Unknown potential type: {potential_type}")
        return torch.exp(-(x - 0.5) ** 2)

    def _jets(self, inputs, predictions):                # :182-200
        u = self.get_complete_solution(inputs, predictions)
        u_x = d_dx(u, inputs)
        return u, u_x, d_dx(u_x, inputs)

    def pde_loss(self, inputs, predictions, gamma, p, potential_type="gaussian", precomputed_potential=None):
        return FinalGPE1D.pde_loss(self, inputs, predictions, gamma, p, potential_type, precomputed_potential)


class VaryPotentialGPE1D(FinalGPE1D):
    """The beta-sweep classes `GrossPitaevskiiPINN(layers, hbar, m, mode, beta, [L], use_residual, use_perturbation)`
    (use_residual is stored and never read):

    family='harmonic'          R/vary_potential_parameter_harmonic.py:51-342 -- base = box sine with L = 5 (:152-160),
                               plain forward (:145-149), V = beta/2 omega^2 (x - c)^2 with omega = 10, c = 2.5 built in
                               compute_potential(x, beta) (:231-246), potential term V u (:296);
    family='box_and_gaussian'  R/vary_potential_parameter_box_and_gaussian.py:52-229 -- base = box sine with L = 1,
                               forward times sin(pi x) (:119-130), V in {0, exp(-x^2/2)} (:141-151), term beta V u (:189);
    family='gravity_well'      R/vary_potential_parameter_gravity_well.py:52-261 -- Airy base with the np.gradient
                               second derivative (:135-173), V = x, term beta V u, mu as sum/sum (:227-229).

    pde_loss(inputs, predictions, gamma, beta, p, potential_type, precomputed_potential) -> (pde_loss, lambda_pde)."""

    _DEFAULT_PT = {"harmonic": "harmonic", "box_and_gaussian": "gaussian", "gravity_well": "gravity_well"}

    def __init__(self, layers, hbar=1.0, m=1.0, mode=0, beta=1.0, L=None, use_residual=True, use_perturbation=True,
                 family="harmonic"):
        L = {"harmonic": 5.0, "box_and_gaussian": 1.0, "gravity_well": 1.0}[family] if L is None else L
        super().__init__(layers, hbar, m, mode, 0.0, use_perturbation,
                         potential="gravity_well" if family == "gravity_well" else "box", L=L)
        self.family, self.beta, self.use_residual = family, beta, use_residual

    def forward(self, inputs):
        out = self.network(inputs)
        if self.family == "box_and_gaussian":
            out = out * torch.sin(torch.pi * inputs)
        return out

    def compute_potential(self, x, *args, **kw):
        if self.family == "harmonic":                    # (x, beta, potential_type="harmonic", omega=10, center=2.5)
            beta = args[0]
            pt = args[1] if len(args) > 1 else kw.get("potential_type", "harmonic")
            if pt != "harmonic":
                raise ValueError(f"Unknown potential type: {pt}")
            return beta * 0.5 * kw.get("omega", 10.0) ** 2 * (x - kw.get("center", 2.5)) ** 2
        pt = args[0] if args else kw.get("potential_type", self._DEFAULT_PT[self.family])
        if self.family == "box_and_gaussian":
            if pt == "box":
                return torch.zeros_like(x)
            if pt == "gaussian":
                return torch.exp((-x ** 2) / 2)
        elif pt == "gravity_well":
            return x.clone()
        raise ValueError(f"Unknown potential type: {pt}")

    def _jets(self, inputs, predictions):
        if self.family == "gravity_well":
            return FinalGPE1D._jets(self, inputs, predictions)
        # the two box-base scripts differentiate base + perturbation by autograd (harmonic :283-299)
        u = self.get_complete_solution(inputs, predictions) if self.use_perturbation else predictions
        u_x = d_dx(u, inputs)
        return u, u_x, d_dx(u_x, inputs)

    def pde_loss(self, inputs, predictions, gamma, beta, p, potential_type=None, precomputed_potential=None):
        pt = potential_type or self._DEFAULT_PT[self.family]
        u, _, u_xx = self._jets(inputs, predictions)
        if precomputed_potential is not None:
            V = precomputed_potential
        else:
            V = self.compute_potential(inputs, beta, pt) if self.family == "harmonic" else self.compute_potential(inputs, pt)
        pot = V * u if self.family == "harmonic" else beta * V * u
        hu = -u_xx + pot + gamma * u ** p
        if self.family == "gravity_well":
            mu = torch.sum(u * hu) / torch.sum(u ** 2)
        else:
            mu = torch.mean(u * hu) / torch.mean(u ** 2)
        return torch.mean((hu - mu * u) ** 2), mu


class PerturbGammaBetaGPE1D(FinalGPE1D):
    """R/test_perturbing_gamma_and_beta.py:53-198 `GrossPitaevskiiPINN(layers, hbar, m, mode, beta, gamma,
    use_perturbation)`: Hermite base, V = self.beta x^2 (:126-135), interaction self.gamma u**p (:171);
    pde_loss(inputs, predictions, p, potential_type, precomputed_potential) (:137-181)."""

    def __init__(self, layers, hbar=1.0, m=1.0, mode=0, beta=1.0, gamma=0.0, use_perturbation=True):
        super().__init__(layers, hbar, m, mode, gamma, use_perturbation, potential="harmonic")
        self.beta = beta

    def compute_potential(self, x, potential_type="harmonic", **kw):
        if potential_type != "harmonic":
            raise ValueError(f"Unknown potential type: {potential_type}")
        return self.beta * (x ** 2)

    def pde_loss(self, inputs, predictions, p, potential_type="harmonic", precomputed_potential=None):
        return FinalGPE1D.pde_loss(self, inputs, predictions, self.gamma, p, potential_type, precomputed_potential)


def variant_train_terms(model, X, boundary_points, boundary_values, dx, pde_args, perturb_const, normal_const):
    """Loop body shared by the beta-sweep scripts (R/vary_potential_parameter_harmonic.py:455-480,
    R/test_perturbing_gamma_and_beta.py:402-429): as final_train_terms, pde_loss(X, u_pred, *pde_args)."""
    u_pred = perturb_const * model.forward(X) / normal_const
    bc = model.boundary_loss(boundary_points, boundary_values)
    nrm = model.normalization_loss(model.get_complete_solution(X, u_pred), dx)
    pde, mu = model.pde_loss(X, u_pred, *pde_args)
    return pde + 10.0 * bc + 20.0 * nrm, pde, bc, nrm, mu


# --------------------------------------------------------------------------------------
# F-opt: S/gross_pitaevskii_1D_Different_Optimizers.py (BASELINE config 2 family)
# --------------------------------------------------------------------------------------
class OptGPE1D(nn.Module):
    """S/gross_pitaevskii_1D_Different_Optimizers.py:27-440."""

    def __init__(self, layers, hbar=1.0, m=1.0, g=100.0, mode=0, beta_init=1.0, alpha_init=1.0, decay_rate=0.001):
        super().__init__()
        self.layers = layers
        self.network = sequential_mlp(layers, nn.Tanh)
        self.g, self.hbar, self.m, self.mode = g, hbar, m, mode
        self.beta, self.alpha, self.decay_rate, self.iteration = beta_init, alpha_init, decay_rate, 0

    def weighted_hermite(self, x, n):
        return hermite_state_constH(x, n)

    def forward(self, inputs):
        return self.network(inputs)

    compute_potential = RootGPE1D.compute_potential

    def get_complete_solution(self, x, perturbation, prev_prediction=None):
        """:175-213 (the norm it computes is unused)."""
        if prev_prediction is None:
            return self.weighted_hermite(x, self.mode) + self.alpha * perturbation
        return self.beta * prev_prediction(x) + self.alpha * perturbation

    def boundary_loss(self, boundary_points, boundary_values, prev_prediction=None):
        u = self.get_complete_solution(boundary_points, self.forward(boundary_points), prev_prediction)
        return torch.mean((u - boundary_values) ** 2)

    def riesz_loss(self, predictions, inputs, eta, potential_type, precomputed_potential=None, prev_prediction=None):
        """:239-282, mean form."""
        u = self.get_complete_solution(inputs, predictions, prev_prediction)
        u_x = d_dx(u, inputs)
        V = precomputed_potential if precomputed_potential is not None else self.compute_potential(inputs, potential_type)
        return 0.5 * (torch.mean(u_x ** 2) + torch.mean(V * u ** 2) + 0.5 * eta * torch.mean(u ** 4))

    def pde_loss(self, inputs, predictions, eta, potential_type, precomputed_potential=None, prev_prediction=None):
        """:284-346: energy-quotient mu, r = -u'' + V u + eta |u^2| u - mu u."""
        u = self.get_complete_solution(inputs, predictions, prev_prediction)
        u_x = d_dx(u, inputs)
        u_xx = d_dx(u_x, inputs)
        V = precomputed_potential if precomputed_potential is not None else self.compute_potential(inputs, potential_type)
        mu = torch.mean(u_x ** 2 + V * u ** 2 + eta * u ** 4) / torch.mean(u ** 2)
        r = -u_xx + V * u + eta * torch.abs(u ** 2) * u - mu * u
        return torch.mean(r ** 2), r, mu

    def symmetry_loss(self, collocation_points, lb, ub):
        """:348-375."""
        a = self.forward(collocation_points)
        b = self.forward((lb + ub) - collocation_points)
        return torch.mean((a - b) ** 2)

    def total_loss(self, collocation_points, boundary_points, boundary_values, eta, lb, ub, weights,
                   potential_type, precomputed_potential=None, prev_prediction=None):
        """:377-440: returns (total, data, riesz, pde, norm); the weighted 5-term sum is computed
        and then replaced by data + pde/len + riesz/len (:438).  `weighted` is exposed as a 6th
        return-less attribute for tests."""
        V = precomputed_potential if precomputed_potential is not None else self.compute_potential(collocation_points, potential_type)
        data = self.boundary_loss(boundary_points, boundary_values, prev_prediction)
        riesz = self.riesz_loss(self.forward(collocation_points), collocation_points, eta, potential_type, V, prev_prediction)
        pde, _, _ = self.pde_loss(collocation_points, self.forward(collocation_points), eta, potential_type, V, prev_prediction)
        norm = (torch.norm(self.forward(collocation_points), p=2) - 1) ** 2
        sym = self.symmetry_loss(collocation_points, lb, ub)
        length = ub - lb
        terms = [data, riesz / length, pde / length, norm, sym]
        self._terms, self._raw = terms, (data, riesz, pde, norm, sym)
        self.last_weighted_sum = sum(w * t for w, t in zip(weights, terms))
        total = data + pde / length + riesz / length
        return total, data, riesz, pde, norm


# --------------------------------------------------------------------------------------
# G2D: S/gross_pitaevskii_2D.py (BASELINE configs 4/5 family) -- keeps the [N]x[N,1] broadcast
# --------------------------------------------------------------------------------------
class _OptAdditive(OptGPE1D):
    """Shared by the two later copies of the F-opt class: ctor (layers, hbar, m, g, mode, prev_prediction=None), no
    alpha / beta, u = base + predictions with base = psi_n or prev_prediction(x)
    (_Self_Adaptive.py:32-61,178-206; _and_Nonlinearity_Powers.py:32-55,172-198)."""

    def __init__(self, layers, hbar=1.0, m=1.0, g=100.0, mode=0, prev_prediction=None):
        super().__init__(layers, hbar, m, g, mode)

    def get_complete_solution(self, x, perturbation, prev_prediction=None):
        base = self.weighted_hermite(x, self.mode) if prev_prediction is None else prev_prediction(x)
        return base + perturbation


class SelfAdaptiveGPE1D(_OptAdditive):
    """S/gross_pitaevskii_1D_Different_Optimizers_Self_Adaptive.py:27-429: five learnable log-weights."""

    def __init__(self, layers, hbar=1.0, m=1.0, g=100.0, mode=0, prev_prediction=None):
        super().__init__(layers, hbar, m, g, mode, prev_prediction)
        self.log_alpha = nn.Parameter(torch.zeros(5, dtype=torch.float32))

    def total_loss(self, collocation_points, boundary_points, boundary_values, eta, lb, ub, weights,
                   potential_type, precomputed_potential=None, prev_prediction=None):
        """:356-429 -> (total, data, riesz, pde, norm, sym, alpha), total = sum_i exp(log_alpha_i) w_i term_i."""
        OptGPE1D.total_loss(self, collocation_points, boundary_points, boundary_values, eta, lb, ub, weights,
                            potential_type, precomputed_potential, prev_prediction)
        alpha = torch.exp(self.log_alpha)
        total = sum(alpha[i] * weights[i] * t for i, t in enumerate(self._terms))
        data, riesz, pde, norm, sym = self._raw
        return total, data, riesz, pde, norm, sym, alpha


class PowersGPE1D(_OptAdditive):
    """S/gross_pitaevskii_1D_Different_Optimizers_and_Nonlinearity_Powers.py:27-407: nonlinearity |u|^(p-1) u."""

    def riesz_loss(self, predictions, inputs, eta, potential_type, precomputed_potential=None, prev_prediction=None, p=2):
        """:200-248."""
        u = self.get_complete_solution(inputs, predictions, prev_prediction)
        u_x = d_dx(u, inputs)
        V = precomputed_potential if precomputed_potential is not None else self.compute_potential(inputs, potential_type)
        return 0.5 * (torch.mean(u_x ** 2) + torch.mean(V * u ** 2) + 2 * eta / (p + 1) * torch.mean(torch.abs(u) ** (p + 1)))

    def pde_loss(self, inputs, predictions, eta, potential_type, precomputed_potential=None, prev_prediction=None, p=2):
        """:250-317."""
        u = self.get_complete_solution(inputs, predictions, prev_prediction)
        u_x = d_dx(u, inputs)
        u_xx = d_dx(u_x, inputs)
        V = precomputed_potential if precomputed_potential is not None else self.compute_potential(inputs, potential_type)
        mu = torch.mean(u_x ** 2 + V * u ** 2 + eta * torch.abs(u) ** (p + 1)) / torch.mean(u ** 2)
        r = -u_xx + V * u + eta * torch.abs(u) ** (p - 1) * u - mu * u
        return torch.mean(r ** 2), r, mu

    def total_loss(self, collocation_points, boundary_points, boundary_values, eta, lb, ub, weights,
                   potential_type, precomputed_potential=None, prev_prediction=None, p=2):
        """:348-407 -> (weighted five-term sum, data, riesz, pde, norm)."""
        V = precomputed_potential if precomputed_potential is not None else self.compute_potential(collocation_points, potential_type)
        data = self.boundary_loss(boundary_points, boundary_values, prev_prediction)
        riesz = self.riesz_loss(self.forward(collocation_points), collocation_points, eta, potential_type, V, prev_prediction, p)
        pde, _, _ = self.pde_loss(collocation_points, self.forward(collocation_points), eta, potential_type, V, prev_prediction, p)
        norm = (torch.norm(self.forward(collocation_points), p=2) - 1) ** 2
        sym = self.symmetry_loss(collocation_points, lb, ub)
        length = ub - lb
        total = sum(w * t for w, t in zip(weights, [data, riesz / length, pde / length, norm, sym]))
        return total, data, riesz, pde, norm


class ModesGPE1D(OptGPE1D):
    """S/gross_pitaevskii_1D_Different_Modes.py:27-450: the mode is an argument of every loss method, the previous
    solution enters without beta (:184-213), total = data + pde/len + riesz/len + norm (:448)."""

    def get_complete_solution(self, x, perturbation, prev_prediction=None, mode=0):
        base = self.weighted_hermite(x, mode) if prev_prediction is None else prev_prediction(x)
        return base + self.alpha * perturbation

    def boundary_loss(self, boundary_points, boundary_values, prev_prediction=None, mode=0):
        u = self.get_complete_solution(boundary_points, self.forward(boundary_points), prev_prediction, mode)
        return torch.mean((u - boundary_values) ** 2)

    def riesz_loss(self, predictions, inputs, eta, potential_type, precomputed_potential=None, prev_prediction=None, mode=0):
        u = self.get_complete_solution(inputs, predictions, prev_prediction, mode)
        u_x = d_dx(u, inputs)
        V = precomputed_potential if precomputed_potential is not None else self.compute_potential(inputs, potential_type)
        return 0.5 * (torch.mean(u_x ** 2) + torch.mean(V * u ** 2) + 0.5 * eta * torch.mean(u ** 4))

    def pde_loss(self, inputs, predictions, eta, potential_type, precomputed_potential=None, prev_prediction=None, mode=0):
        u = self.get_complete_solution(inputs, predictions, prev_prediction, mode)
        u_x = d_dx(u, inputs)
        u_xx = d_dx(u_x, inputs)
        V = precomputed_potential if precomputed_potential is not None else self.compute_potential(inputs, potential_type)
        mu = torch.mean(u_x ** 2 + V * u ** 2 + eta * u ** 4) / torch.mean(u ** 2)
        r = -u_xx + V * u + eta * torch.abs(u ** 2) * u - mu * u
        return torch.mean(r ** 2), r, mu

    def total_loss(self, collocation_points, boundary_points, boundary_values, eta, lb, ub, weights,
                   potential_type, precomputed_potential=None, prev_prediction=None, mode=0):
        """:383-450."""
        V = precomputed_potential if precomputed_potential is not None else self.compute_potential(collocation_points, potential_type)
        data = self.boundary_loss(boundary_points, boundary_values, prev_prediction, mode)
        riesz = self.riesz_loss(self.forward(collocation_points), collocation_points, eta, potential_type, V, prev_prediction, mode)
        pde, _, _ = self.pde_loss(collocation_points, self.forward(collocation_points), eta, potential_type, V, prev_prediction, mode)
        norm = (torch.norm(self.forward(collocation_points), p=2) - 1) ** 2
        length = ub - lb
        return data + pde / length + riesz / length + norm, data, riesz, pde, norm


class GPE2D(nn.Module):
    """S/gross_pitaevskii_2D.py:16-274.  `potential` extends compute_potential beyond the
    reference's Gaussian so the BASELINE harmonic / optical-lattice configs can be stated."""

    def __init__(self, layers, hbar=1.0, m=1.0, g=100.0, potential="gaussian"):
        super().__init__()
        self.layers = layers
        self.network = sequential_mlp(layers, nn.Tanh)
        self.hbar, self.m, self.g, self.potential = hbar, m, g, potential

    def forward(self, x):
        return self.network(x)

    def boundary_loss(self, x_bc, y_bc):
        """:83-109: target forced to zero, x10."""
        return torch.mean(self.forward(x_bc) ** 2) * 10

    def compute_potential(self, inputs, V0=1.0, x0=np.pi / 2, y0=np.pi / 2, sigma=0.5):
        """:244-274 -> shape [N]."""
        x, y = inputs[:, 0], inputs[:, 1]
        if self.potential == "gaussian":
            return V0 * torch.exp(-((x - x0) ** 2 + (y - y0) ** 2) / (2 * sigma ** 2))
        if self.potential == "harmonic":
            return 0.5 * (x ** 2 + y ** 2)
        if self.potential == "lattice":                  # V0 cos^2(kx) form of NBroot cell 6:73-76 per axis + trap
            k = 2 * np.pi / 5.0
            return V0 * torch.cos(k * x) ** 2 + V0 * torch.cos(k * y) ** 2 + 0.5 * (x ** 2 + y ** 2)
        raise ValueError(self.potential)

    def riesz_loss(self, predictions, inputs):
        """:112-151, sum form (V unsqueezed -> no broadcast bug here)."""
        u = predictions
        g_ = d_dx(u, inputs)
        V = self.compute_potential(inputs).unsqueeze(1)
        return 0.5 * (torch.sum(g_ ** 2) + torch.sum(V * u ** 2) + 0.5 * self.g * torch.sum(u ** 4))

    def pde_loss(self, inputs, predictions):
        """:154-213.  u is [N,1]; u_x, u_xx, V are [N]  => lambda and the residual broadcast to
        [N,N] exactly as in the reference."""
        u = predictions
        u_x = d_dx(u, inputs)[:, 0]
        u_y = d_dx(u, inputs)[:, 1]
        u_xx = d_dx(u_x, inputs)[:, 0]
        u_yy = d_dx(u_y, inputs)[:, 1]
        lap = u_xx + u_yy
        V = self.compute_potential(inputs)
        mu = torch.mean(u_x ** 2 + u_y ** 2 + V * u ** 2 + self.g * u ** 4) / torch.mean(u ** 2)
        r = -lap + V * u + self.g * torch.abs(u ** 2) * u - (mu * u)
        L_f = 1 / (torch.mean(u ** 2) + 1e-2)
        L_lambda = 1 / (mu ** 2 + 1e-6)
        return torch.mean(r ** 2) + L_f + L_lambda, r, mu

    def loss(self, x, x_bc, u_bc):
        """:215-242."""
        data = self.boundary_loss(x_bc, u_bc)
        riesz = self.riesz_loss(self.forward(x), x)
        pde, _, _ = self.pde_loss(x, self.forward(x))
        return data + pde + riesz

    total_loss = loss            # S/gross_pitaevskii_2D_minimal.py:201-222: the same sum under another name


class ReLoBRaLoGPE2D(nn.Module):
    """S/gross_pitaevskii_2D_ReLoBRaLo.py:13-341.  The file keeps the 1-D formulas on [N,2] inputs: V = exp(-(x-a)^2)
    is [N,2] (:94-126), u_x = grad u [N,2], u_xx = grad(sum u_x) = Hessian . (1,1) [N,2] (:200-201), so the residual is
    [N,2] with a shared scalar lambda; the mirror is y -> -y (:232-257), the norm term (||NN||_2 - 1)^2 (:285) and the
    five terms are weighted by ReLoBRaLo's .item()-based softmax (:294-339)."""

    def __init__(self, layers, hbar=1.0, m=1.0, g=100.0, alpha=0.999, temperature=1.0, rho=0.9999):
        super().__init__()
        self.layers = layers
        self.network = sequential_mlp(layers, nn.Tanh)
        self.g, self.hbar, self.m = g, hbar, m
        self.alpha, self.temperature, self.rho = alpha, temperature, rho
        self.call_count = 0
        self.lambdas, self.last_losses, self.init_losses = [1.0] * 5, [1.0] * 5, [1.0] * 5

    def forward(self, inputs):
        return self.network(inputs)

    def compute_potential(self, x, potential_type="gaussian", a=0.5, l=1.0):
        if potential_type == "gaussian":
            return torch.exp(-(x - a) ** 2)
        return torch.sin(torch.pi * (x - (a - l / 2)) / l)

    def boundary_loss(self, boundary_points, boundary_values):
        return torch.mean((self.forward(boundary_points) - boundary_values) ** 2)

    def riesz_loss(self, predictions, inputs, eta):
        """:147-180, mean form."""
        u = predictions
        u_x = d_dx(u, inputs)
        V = self.compute_potential(inputs)
        return 0.5 * (torch.mean(u_x ** 2) + torch.mean(V * u ** 2) + 0.5 * eta * torch.mean(u ** 4))

    def pde_loss(self, inputs, predictions, eta):
        """:182-230 -> (pde_loss, pde_residual [N,2], lambda_pde)."""
        u = predictions
        u_x = d_dx(u, inputs)
        u_xx = d_dx(u_x, inputs)
        V = self.compute_potential(inputs)
        lam = torch.mean(u_x ** 2 + V * u ** 2 + eta * u ** 4) / torch.mean(u ** 2)
        r = -u_xx + V * u + eta * torch.abs(u ** 2) * u - lam * u
        return torch.mean(r ** 2), r, lam

    def symmetry_loss(self, collocation_points):
        refl = collocation_points.clone()
        refl[:, 1] = -refl[:, 1]
        return torch.mean((self.forward(collocation_points) - self.forward(refl)) ** 2)

    def relobralo_weights(self, losses):
        """:294-335: the weight update from the terms' host values (one torch.bernoulli draw per call)."""
        vals = [t.item() for t in losses]
        if self.call_count == 0:
            self.lambdas, self.last_losses, self.init_losses = [1.0] * len(vals), list(vals), list(vals)
        hat = [vals[i] / (self.last_losses[i] * self.temperature + 1e-8) for i in range(len(vals))]
        hat = torch.softmax(torch.tensor(hat) - max(hat), dim=0).tolist()
        hat0 = [vals[i] / (self.init_losses[i] * self.temperature + 1e-8) for i in range(len(vals))]
        hat0 = torch.softmax(torch.tensor(hat0) - max(hat0), dim=0).tolist()
        rho = torch.bernoulli(torch.tensor(self.rho))
        alpha = self.alpha if self.call_count > 1 else (0.0 if self.call_count == 1 else 1.0)
        self.lambdas = [float(rho * alpha * self.lambdas[i] + (1 - rho) * alpha * hat0[i] + (1 - alpha) * hat[i])
                        for i in range(len(vals))]
        self.last_losses = list(vals)
        self.call_count += 1
        return [lam * w for lam, w in zip(self.lambdas, [500.0, 1.0, 2.0, 100.0, 500])]

    def total_loss(self, collocation_points, boundary_points, boundary_values, eta):
        """:259-341 -> (total, data, riesz, pde, norm)."""
        data = self.boundary_loss(boundary_points, boundary_values)
        riesz = self.riesz_loss(self.forward(collocation_points), collocation_points, eta)
        pde, _, _ = self.pde_loss(collocation_points, self.forward(collocation_points), eta)
        norm = (torch.norm(self.forward(collocation_points), p=2) - 1) ** 2
        sym = self.symmetry_loss(collocation_points)
        losses = [data, riesz, pde, norm, sym]
        total = sum(w * t for w, t in zip(self.relobralo_weights(losses), losses))
        return total, data, riesz, pde, norm


def gpe2d_pointwise_terms(model, x, x_bc):
    """The *intended* (pointwise, all [N,1]) mathematics of S/gross_pitaevskii_2D.py:154-242:
    returns (total, bc, pde_with_regs, riesz, mu)."""
    u = model.forward(x)
    gr = d_dx(u, x)
    u_x, u_y = gr[:, [0]], gr[:, [1]]
    lap = d_dx(u_x, x)[:, [0]] + d_dx(u_y, x)[:, [1]]
    V = model.compute_potential(x).unsqueeze(1)
    mu = torch.mean(u_x ** 2 + u_y ** 2 + V * u ** 2 + model.g * u ** 4) / torch.mean(u ** 2)
    r = -lap + V * u + model.g * torch.abs(u ** 2) * u - mu * u
    pde = torch.mean(r ** 2) + 1 / (torch.mean(u ** 2) + 1e-2) + 1 / (mu ** 2 + 1e-6)
    riesz = 0.5 * (torch.sum(gr ** 2) + torch.sum(V * u ** 2) + 0.5 * model.g * torch.sum(u ** 4))
    bc = model.boundary_loss(x_bc, None)
    return bc + pde + riesz, bc, pde, riesz, mu


# --------------------------------------------------------------------------------------
# H-sq: S/helmholtz_2D.py (BASELINE config 3)
# --------------------------------------------------------------------------------------
class Helmholtz2D(nn.Module):
    """S/helmholtz_2D.py:19-132.  `activation='leaky_relu'` is the reference default (:31);
    'tanh' follows S/helmholtz_2D_circle_3.py:31.  a1,a2 generalise the source to
    sin(a1 pi x) sin(a2 pi y) (BASELINE config 3); a1=a2=1 is the reference.  lb / ub: the min-max feature
    scaling of S/helmholtz_2D_circle.py:89-94 (module globals there)."""

    def __init__(self, layers, activation="leaky_relu", a1=1, a2=1, lb=None, ub=None):
        super().__init__()
        self.lb, self.ub = lb, ub
        self.activation = nn.LeakyReLU() if activation == "leaky_relu" else nn.Tanh()
        self.loss_function = nn.MSELoss()
        self.linears = nn.ModuleList([nn.Linear(layers[i], layers[i + 1]) for i in range(len(layers) - 1)])
        self.a1, self.a2 = a1, a2
        for lin in self.linears:
            nn.init.xavier_normal_(lin.weight)
            nn.init.zeros_(lin.bias)

    def forward(self, x):
        if self.lb is not None:                                   # helmholtz_2D_circle.py:89-94
            l_b = torch.as_tensor(self.lb, dtype=x.dtype, device=x.device)
            u_b = torch.as_tensor(self.ub, dtype=x.dtype, device=x.device)
            x = (x - l_b) / (u_b - l_b)
        for lin in self.linears[:-1]:
            x = self.activation(lin(x))
        return self.linears[-1](x)

    def loss_pde(self, x_f, k):
        """:58-89 ([N] + [N,1] -> [N,N] residual, MSE over all N^2 entries)."""
        g = x_f.clone().requires_grad_(True)
        u = self.forward(g)
        u_x = d_dx(u, g)[:, 0]
        u_y = d_dx(u, g)[:, 1]
        lap = d_dx(u_x, g)[:, 0] + d_dx(u_y, g)[:, 1]
        q = k ** 2 * torch.sin(self.a1 * np.pi * g[:, [0]]) * torch.sin(self.a2 * np.pi * g[:, [1]])
        f = lap + k ** 2 * u - q
        return self.loss_function(f, torch.zeros_like(f))

    def loss_pde_pointwise(self, x_f, k):
        """Intended mathematics: everything [N,1]."""
        g = x_f.clone().requires_grad_(True)
        u = self.forward(g)
        gr = d_dx(u, g)
        lap = d_dx(gr[:, [0]], g)[:, [0]] + d_dx(gr[:, [1]], g)[:, [1]]
        q = k ** 2 * torch.sin(self.a1 * np.pi * g[:, [0]]) * torch.sin(self.a2 * np.pi * g[:, [1]])
        return torch.mean((lap + k ** 2 * u - q) ** 2)

    def loss_data(self, x_u, u_true):
        return self.loss_function(self.forward(x_u), u_true)

    def compute_loss(self, x_f, x_u, u_true, k):
        return self.loss_pde(x_f, k) + self.loss_data(x_u, u_true)


class HelmholtzCircle(Helmholtz2D):
    """`HelmholtzPINN` of S/helmholtz_2D_circle.py:28-172: tanh MLP (:53) behind the min-max feature scaling of
    :89-94, Helmholtz residual on the disc written in polar form from Cartesian derivatives.  The shapes are the
    reference's: coordinates and u are [N,1], the autograd columns are [N], so every product of the two is an
    [N,N] table (row = the point the coordinate belongs to, column = the point the derivative belongs to) and the
    MSE runs over all N^2 entries."""

    def __init__(self, layers, lb=(-1.0, -1.0), ub=(1.0, 1.0)):
        super().__init__(layers, activation="tanh", lb=list(lb), ub=list(ub))

    def loss_PDE(self, x_f, k=None):
        """:112-172.  The argument k is ignored by the reference, which recomputes k = pi sqrt(m^2 + n^2), m = n = 1."""
        from scipy.special import jn
        X = x_f.clone().requires_grad_(True)
        xc, yc = X[:, [0]], X[:, [1]]                                    # [N,1]
        r = torch.clamp(torch.sqrt(xc ** 2 + yc ** 2), min=1e-8)
        theta = torch.atan2(yc, xc)
        u = self.forward(X)                                              # [N,1]
        du = d_dx(u, X)                                                  # [N,2]
        ux, uy = du[:, 0], du[:, 1]                                      # [N]
        d2 = d_dx(du, X)                    # grad of sum(u_x + u_y): columns u_xx + u_xy, u_xy + u_yy   (:137-140)
        s0, s1 = d2[:, 0], d2[:, 1]
        u_r = (xc / r) * ux + (yc / r) * uy                              # [N,N]
        u_rr = (xc ** 2 / r ** 2) * s0 + (2 * xc * yc / r ** 2) * s1 + (yc ** 2 / r ** 2) * s1   # (:144-145) as written
        u_th = -yc * ux + xc * uy                                        # [N,N]
        dth = d_dx(u_th, X)                                              # sum over the whole table  (:149-150)
        u_thth = -yc * dth[:, 0] + xc * dth[:, 1]                        # [N,N]
        n_ang, m_rad = 1, 1
        kk = math.sqrt(math.pi ** 2 * (m_rad ** 2 + n_ang ** 2))
        # the reference builds k as a float32 tensor and J_n as float64 values of the float32 product k r  (:156-159)
        k32 = torch.sqrt(torch.tensor(np.pi ** 2 * (m_rad ** 2 + n_ang ** 2), device=x_f.device))
        kr = k32.cpu().numpy() * r.detach().cpu().numpy()
        bessel = torch.tensor(jn(n_ang, kr), dtype=torch.float64, device=x_f.device)
        kt = k32 if x_f.dtype == torch.float32 else torch.tensor(kk, dtype=x_f.dtype, device=x_f.device)
        q = kt ** 2 * bessel * torch.cos(n_ang * theta)                  # [N,1] float64
        f = u_rr + (1 / r) * u_r + (1 / r ** 2) * u_thth + kt ** 2 * u - q
        return self.loss_function(f.view(-1), torch.zeros_like(f).view(-1))


class HelmholtzCircle2(Helmholtz2D):
    """`HelmholtzPINN` of S/helmholtz_2D_circle_2.py:28-179: LeakyReLU MLP (:53) behind the min-max feature scaling
    (:89-94); the inputs ARE the polar coordinates (r, theta) and the residual is written directly in them.  Shapes as
    in the reference: r, theta are [N] and the autograd columns [N,1], so the residual is an [N,N] table."""

    def __init__(self, layers, lb=(0.0, 0.0), ub=(1.0, 2 * math.pi)):
        super().__init__(layers, activation="leaky_relu", lb=list(lb), ub=list(ub))

    def loss_PDE(self, x_f, n=1, m=1):
        """:108-160."""
        g = x_f.clone().requires_grad_(True)
        r, theta = g[:, 0], g[:, 1]
        u = self.forward(g)
        du = d_dx(u, g)
        u_r, u_t = du[:, [0]], du[:, [1]]
        u_rr = d_dx(u_r, g)[:, [0]]
        u_tt = d_dx(u_t, g)[:, [1]]
        if x_f.dtype == torch.float32:
            k = torch.sqrt(torch.tensor(np.pi ** 2 * (m ** 2 + n ** 2), dtype=torch.float32, device=g.device))
        else:
            k = torch.tensor(math.sqrt(math.pi ** 2 * (m ** 2 + n ** 2)), dtype=x_f.dtype, device=g.device)
        q = (k ** 2) * torch.sin(n * r * theta) * torch.sin(m * r * theta)
        f = u_rr + (1 / r) * u_r + (1 / r ** 2) * u_tt + k ** 2 * u - q
        return self.loss_function(f, torch.zeros_like(f))


class HelmholtzCircle3(Helmholtz2D):
    """`HelmholtzPINN` of S/helmholtz_2D_circle_3.py:20-132: tanh MLP, no input scaling, polar residual from Cartesian
    autograd columns ([N]) against u [N,1] -> [N,N]; u_thetatheta = d(u_theta)/d(theta) with u_x, u_y held fixed
    (theta is not a leaf: the graph reaches it only through cos / sin) divided by r (:96-97); source k^2 J_0(k r) cos(n theta)."""

    def __init__(self, layers):
        super().__init__(layers, activation="tanh")

    def loss_pde(self, x_f, n, k):
        """:58-110."""
        g = x_f.clone().requires_grad_(True)
        u = self.forward(g)
        r = torch.sqrt(g[:, 0] ** 2 + g[:, 1] ** 2)
        theta = torch.atan2(g[:, 1], g[:, 0])
        du = d_dx(u, g)
        u_x, u_y = du[:, 0], du[:, 1]
        u_r = u_x * torch.cos(theta) + u_y * torch.sin(theta)
        u_theta = -u_x * torch.sin(theta) + u_y * torch.cos(theta)
        u_xx = d_dx(u_x, g)[:, 0]
        u_yy = d_dx(u_y, g)[:, 1]
        u_rr = u_xx * torch.cos(theta) ** 2 + u_yy * torch.sin(theta) ** 2
        u_tt = torch.autograd.grad(u_theta, theta, grad_outputs=torch.ones_like(u_theta), create_graph=True)[0] / r
        lap = u_rr + (1 / r) * u_r + (1 / r ** 2) * u_tt
        q = k ** 2 * torch.special.bessel_j0(k * r) * torch.cos(n * theta)
        res = lap + k ** 2 * u - q
        return self.loss_function(res, torch.zeros_like(res))


# --------------------------------------------------------------------------------------
# helpers shared by tests / bench
# --------------------------------------------------------------------------------------
def linear_layers(module):
    """The nn.Linear modules of an oracle (or reference) model, input to output."""
    return [m for m in module.modules() if isinstance(m, nn.Linear)]


def param_grads(model, scalar):
    """d scalar / d theta as a list ordered like `model.parameters()` (zeros for unused)."""
    ps = list(model.parameters())
    gs = torch.autograd.grad(scalar, ps, allow_unused=True)
    return [torch.zeros_like(p) if g is None else g for p, g in zip(ps, gs)]


class LearnableKHelmholtz2D(nn.Module):
    """`PINN(layers, ub, lb)` of Notebooks/Helmholtz/Helmholtz_learnable_k_and_bc.ipynb cell 3 (SURVEY #18, 8f-4):
    LeakyReLU MLP held in `layers` (Linear / LeakyReLU alternating), min-max input scaling (:68-73), learnable wave
    number `k` and boundary weight `adaptive_bc_scale` (:35-36).  The notebook checkpoints its Linear layers (:76-80),
    which changes memory, not numbers.  pde_loss (:129-162): `u_xx + u_yy` is [N], `predictions` [N,1], so the residual
    is the [N,N] matrix lap_j - k^2 u_i; loss (:164-192) multiplies the boundary term by the scale a second time."""

    def __init__(self, layers, ub, lb):
        super().__init__()
        self.layers = nn.ModuleList()
        self.ub = torch.tensor(ub, dtype=torch.float32)
        self.lb = torch.tensor(lb, dtype=torch.float32)
        self.adaptive_bc_scale = nn.Parameter(torch.tensor(10.0))
        self.k = nn.Parameter(torch.tensor(np.pi, dtype=torch.float32))
        for i in range(len(layers) - 1):
            self.layers.append(nn.Linear(layers[i], layers[i + 1]))
            if i < len(layers) - 2:
                self.layers.append(nn.LeakyReLU())
        self.activation = nn.LeakyReLU()
        for m in self.layers:
            if isinstance(m, nn.Linear):
                nn.init.xavier_normal_(m.weight)

    def forward(self, x):
        lb, ub = self.lb.view(1, -1).to(x.device), self.ub.view(1, -1).to(x.device)
        x = (x - lb) / (ub - lb)
        for layer in self.layers:
            x = layer(x)
        return x

    def loss_BC(self, x_bc, y_bc):
        return self.adaptive_bc_scale * torch.mean((self.forward(x_bc) - y_bc) ** 2)

    def riesz_loss(self, predictions, inputs):
        g = torch.autograd.grad(predictions, inputs, torch.ones_like(predictions), create_graph=True, retain_graph=True)[0]
        return torch.sum(g ** 2)

    def pde_loss(self, inputs, predictions):
        u_x = torch.autograd.grad(predictions, inputs, torch.ones_like(predictions), create_graph=True)[0]
        u_xx = torch.autograd.grad(u_x[:, 0], inputs, torch.ones_like(u_x[:, 0]), create_graph=True)[0][:, 0]
        u_yy = torch.autograd.grad(u_x[:, 1], inputs, torch.ones_like(u_x[:, 1]), create_graph=True)[0][:, 1]
        return torch.mean((u_xx + u_yy - self.k ** 2 * predictions) ** 2)

    def loss(self, x_bc, y_bc, x_to_train_f):
        loss_u = self.adaptive_bc_scale * self.loss_BC(x_bc, y_bc)
        predictions = self.forward(x_to_train_f)
        return loss_u + self.pde_loss(x_to_train_f, predictions) + self.riesz_loss(predictions, x_to_train_f)
