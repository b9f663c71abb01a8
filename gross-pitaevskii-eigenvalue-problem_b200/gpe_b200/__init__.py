"""gpe_b200 -- B200 (sm_100a) fused PINN train step for the Gross-Pitaevskii / Helmholtz PINNs of
LevBahn/Gross-Pitaevskii-Eigenvalue-problem.  CUDA only: importing without the built
libgpe_b200.so / libgpe_b200_torch.so raises, and there is no CPU fallback."""
from . import _lib, ops  # noqa: F401
from . import distributed, graphs, losses, models, training  # noqa: F401
from ._lib import version  # noqa: F401
from .models import (BoxToGaussianGPE1D, FinalGPE1D, GPE2D, Helmholtz2D, LearnableKHelmholtz2D, ModesGPE1D, OptGPE1D, PaperGPE1D, PerturbGammaBetaGPE1D,  # noqa: F401
                     PowersGPE1D, ReLoBRaLoGPE2D, ResidualBlock, RootGPE1D, SelfAdaptiveGPE1D, ShiftedTanh, VaryPotentialGPE1D, patch)

__all__ = ["ops", "losses", "models", "distributed", "graphs", "training", "version", "RootGPE1D", "FinalGPE1D", "OptGPE1D", "GPE2D",
           "Helmholtz2D", "ShiftedTanh", "VaryPotentialGPE1D", "PerturbGammaBetaGPE1D",
           "PaperGPE1D", "SelfAdaptiveGPE1D", "PowersGPE1D", "ModesGPE1D", "ReLoBRaLoGPE2D",
           "BoxToGaussianGPE1D", "ResidualBlock", "LearnableKHelmholtz2D", "patch"]
