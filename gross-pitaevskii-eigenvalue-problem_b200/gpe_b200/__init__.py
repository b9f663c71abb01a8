"""gpe_b200 -- B200 (sm_100a) fused PINN train step for the Gross-Pitaevskii / Helmholtz PINNs of
LevBahn/Gross-Pitaevskii-Eigenvalue-problem.  CUDA only: importing without the built
libgpe_b200.so raises, and there is no CPU fallback."""
from . import _lib, ops  # noqa: F401
from ._lib import version  # noqa: F401

__all__ = ["ops", "version"]
