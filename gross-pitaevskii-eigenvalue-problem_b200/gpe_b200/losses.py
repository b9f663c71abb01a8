"""Loss terms of the reference PINNs as smooth functions of the fp64 moment vector.

`ops.gram_moments` reduces a shard of collocation points to m[32]: the Gram matrix of the
per-point features f = [1, u, lap u, V, V u, nl(u), q] plus sum|grad u|^2, sum N^2 and the
symmetry sum (layout: include/gpe_b200.h).  Everything below is tiny fp64 torch arithmetic on that
vector, so autograd produces c = dL/dm exactly -- including the dependence of the residual on the
global scalar mu -- and `gpe::gram_adjoints` turns c into per-point adjoints.
Because the sums are fp64 and the per-point products are exact in fp64, forming mean(r^2) from
second moments does not suffer the fp32 cancellation SURVEY.md fact 7 warns about.

With a process group, the moment vector is all-reduced (SUM) first; m[(1,1)] is then the global
point count, so every mean below is a global mean.
"""
import os

import torch

from ._lib import (F_LAP, F_NL, F_ONE, F_Q, F_U, F_V, F_VU, MOM_GRAD2, MOM_NN2, MOM_SYM2, NFEAT, mom)


def sqrt(v):
    """Square root of a float64 torch scalar or of a recorded scalar (expr.Expr)."""
    return torch.sqrt(v) if isinstance(v, torch.Tensor) else v ** 0.5


def fused_terms(moments, build, key=None):
    """The terms `build(*M)` of one loss method as float64 0-dim tensors.  `moments`: Moments objects (one per moment
    vector the terms read).  On CUDA the arithmetic is recorded once and runs as ONE kernel that also returns the
    Jacobian with respect to the moments (expr.py, `gpe::scalar_program`) -- GPE_FUSED_TERMS=0 keeps the float64 torch
    scalars (one small kernel per operator), which is also what CPU moment vectors (the host-logic tests) get."""
    if moments[0].m.is_cuda and os.environ.get("GPE_FUSED_TERMS", "1") != "0":
        from . import expr
        try:
            return expr.evaluate([M.m for M in moments], build, key)
        except TypeError:
            pass        # a coefficient arrived as a tensor (e.g. a learnable wave number): torch scalars carry its gradient
    return tuple(build(*moments))


class Moments:
    """View of the moment vector with named access.  `P(i, j)` = sum_n f_i f_j."""

    def __init__(self, m: torch.Tensor):
        self.m = m

    def P(self, i, j):
        return self.m[mom(i, j)]

    @property
    def n(self):
        return self.m[mom(F_ONE, F_ONE)]

    def mean(self, i, j=F_ONE):
        return self.P(i, j) / self.n

    @property
    def grad2(self):
        return self.m[MOM_GRAD2]

    @property
    def nn2(self):
        return self.m[MOM_NN2]

    @property
    def sym2(self):
        return self.m[MOM_SYM2]

    def quad(self, w):
        """sum_n (w . f_n)^2 for per-feature weights w (dict feature -> scalar / 0-dim tensor)."""
        items = list(w.items())
        total = 0.0
        for a, (i, wi) in enumerate(items):
            for (j, wj) in items[a:]:
                total = total + (1.0 if i == j else 2.0) * wi * wj * self.P(i, j)
        return total

    def dot_u(self, w):
        """sum_n u_n (w . f_n)."""
        return sum(wi * self.P(F_U, i) for i, wi in w.items())


def hamiltonian_weights(kinetic, eta):
    """H u = kinetic * lap u + V u + eta * nl(u) as feature weights
    (kinetic = -1/2: NBroot cell 6:113; -1: R/harmonic_pinn_simulation.py:181)."""
    return {F_LAP: kinetic, F_VU: 1.0, F_NL: eta}


def rayleigh_mu(M: Moments, wH):
    """mu = mean(u Hu) / mean(u^2)  (NBroot cell 6:117-119); the sum/sum form of
    R/gravity_well_pinn_simulation.py:226-228 is the same number."""
    return M.dot_u(wH) / M.P(F_U, F_U)


def energy_mu(M: Moments, eta):
    """mu = mean(|grad u|^2 + V u^2 + eta u nl(u)) / mean(u^2)
    (S/gross_pitaevskii_1D_Different_Optimizers.py:323, pointwise form of S/gross_pitaevskii_2D.py:192)."""
    return (M.grad2 + M.P(F_U, F_VU) + eta * M.P(F_U, F_NL)) / M.P(F_U, F_U)


def residual_mse(M: Moments, wH, mu):
    """mean((Hu - mu u)^2) over the collocation set."""
    w = dict(wH)
    w[F_U] = w.get(F_U, 0.0) - mu
    return M.quad(w) / M.n


def g2d_reference_mu(M: Moments, g):
    """S/gross_pitaevskii_2D.py:192 with its shapes: u_x^2 [N] + V [N] * u^2 [N,1] + g u^4 [N,1]
    broadcasts to [N,N]; its mean is <|grad u|^2> + <V><u^2> + g <u^4>  (SURVEY.md App. C.1)."""
    return (M.grad2 / M.n + M.mean(F_V) * M.mean(F_U, F_U) + g * M.mean(F_U, F_NL)) / M.mean(F_U, F_U)


def g2d_reference_residual_mse(M: Moments, g, mu):
    """mean over the [N,N] matrix r_ij = a_j + V_j u_i + b_i, a = -lap u, b = g nl - mu u
    (S/gross_pitaevskii_2D.py:195-198)."""
    a2 = M.mean(F_LAP, F_LAP)
    wb = {F_NL: g, F_U: -mu}
    b2 = M.quad(wb) / M.n
    b1 = g * M.mean(F_NL) - mu * M.mean(F_U)
    ub = g * M.mean(F_U, F_NL) - mu * M.mean(F_U, F_U)
    return (a2 + M.mean(F_V, F_V) * M.mean(F_U, F_U) + b2 - 2.0 * M.mean(F_LAP, F_V) * M.mean(F_U)
            - 2.0 * M.mean(F_LAP) * b1 + 2.0 * M.mean(F_V) * ub)


def helmholtz_reference_mse(M: Moments, k2):
    """S/helmholtz_2D.py:84-89 with its shapes: f_ij = lap_j + (k^2 u_i - q_i);
    mean f^2 = <lap^2> + 2 <lap><b> + <b^2>, b = k^2 u - q (q already carries its k^2)."""
    wb = {F_U: k2, F_Q: -1.0}
    b1 = k2 * M.mean(F_U) - M.mean(F_Q)
    return M.mean(F_LAP, F_LAP) + 2.0 * M.mean(F_LAP) * b1 + M.quad(wb) / M.n


def helmholtz_pointwise_mse(M: Moments, k2):
    """mean((lap u + k^2 u - q)^2) -- the intended mathematics."""
    return M.quad({F_LAP: 1.0, F_U: k2, F_Q: -1.0}) / M.n


def riesz_energy(M: Moments, eta, reduction="mean", nl_coef=0.5):
    """1/2 (R|grad u|^2 + R V u^2 + nl_coef eta R u nl(u)), R = mean
    (S/gross_pitaevskii_1D_Different_Optimizers.py:278-280) or sum (S/gross_pitaevskii_2D.py:146-149); nl_coef = 1/2
    for u^4, 2/(p+1) for |u|^(p+1) (S/gross_pitaevskii_1D_Different_Optimizers_and_Nonlinearity_Powers.py:244)."""
    s = M.grad2 + M.P(F_U, F_VU) + nl_coef * eta * M.P(F_U, F_NL)
    return 0.5 * (s / M.n if reduction == "mean" else s)
