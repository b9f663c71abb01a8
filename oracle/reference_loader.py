"""Import the REAL reference classes from /root/reference (build container only).

TEST INFRASTRUCTURE (see oracle/__init__.py).  /root/reference does not exist on the GPU box,
so nothing here runs in `-m gpu` tests, smoke() or bench.py; it is used by
tests/golden/make_golden.py to generate fixtures and by CPU tests that skip when the tree is
absent.  No reference source is copied: modules are executed from where they lie.

The scripts import plotting / sampling / optimizer packages that are not installed here
(matplotlib, pyDOE, pytorch_optimizer, adabelief_pytorch, distributed_shampoo); none is on the
arithmetic path, so they are replaced by inert stubs before import.
"""
import importlib.util
import json
import math
import os
import sys
import types

import numpy as np
import torch
import torch.nn as nn

REF_ROOT = "/root/reference"
SRC = os.path.join(REF_ROOT, "Gross-Pitaevskii", "src")
REFINE = os.path.join(SRC, "final", "refine")


def available():
    return os.path.isdir(SRC)


class _Anything:
    """Absorbs any attribute access / call (for matplotlib & friends)."""

    def __getattr__(self, name):
        return _Anything()

    def __call__(self, *a, **k):
        return _Anything()

    def __iter__(self):
        return iter(())

    def update(self, *a, **k):
        pass

    def __setitem__(self, k, v):
        pass

    def __getitem__(self, k):
        return _Anything()


def _stub(name, **attrs):
    mod = types.ModuleType(name)
    mod.__dict__.update(attrs)
    mod.__getattr__ = lambda attr: _Anything()  # PEP 562
    sys.modules[name] = mod
    return mod


def install_stubs():
    from scipy.stats import qmc
    if "matplotlib" not in sys.modules:
        mpl = _stub("matplotlib", use=lambda *a, **k: None, rcParams=_Anything())
        mpl.pyplot = _stub("matplotlib.pyplot", rcParams=_Anything())
        for sub in ("ticker", "gridspec", "cm", "colors", "patches", "lines"):
            setattr(mpl, sub, _stub("matplotlib." + sub))
        _stub("mpl_toolkits")
        _stub("mpl_toolkits.mplot3d")
        _stub("mpl_toolkits.axes_grid1")
    if "pyDOE" not in sys.modules:
        _stub("pyDOE", lhs=lambda d, n=None, samples=None: qmc.LatinHypercube(d=d, seed=1234).random(n or samples))
    for name in ("pytorch_optimizer", "adabelief_pytorch", "distributed_shampoo",
                 "distributed_shampoo.distributed_shampoo", "distributed_shampoo.shampoo_types", "seaborn", "tqdm"):
        if name not in sys.modules:
            _stub(name)


def load_script(relpath, modname=None):
    """Execute a reference .py (its __main__ block is guarded) and return the module."""
    install_stubs()
    path = os.path.join(SRC, relpath)
    modname = modname or "ref_" + os.path.basename(relpath)[:-3]
    spec = importlib.util.spec_from_file_location(modname, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def load_class_by_lines(relpath, first, last, extra_ns=None):
    """For scripts with unguarded module-level loops: exec only a line range."""
    install_stubs()
    from scipy.special import hermite
    from torch.autograd import grad
    with open(os.path.join(SRC, relpath)) as fh:
        lines = fh.readlines()
    ns = dict(torch=torch, nn=nn, np=np, math=math, hermite=hermite, grad=grad,
              device=torch.device("cpu"))
    ns.update(extra_ns or {})
    exec(compile("".join(lines[first - 1:last]), relpath, "exec"), ns)
    return ns


def load_root_notebook_class():
    """`GrossPitaevskiiPINN` of Gross_Pitaevskii_1D_power_Test.ipynb cell 6."""
    from scipy.special import hermite
    with open(os.path.join(REF_ROOT, "Gross_Pitaevskii_1D_power_Test.ipynb")) as fh:
        nb = json.load(fh)
    ns = dict(torch=torch, nn=nn, np=np, math=math, hermite=hermite, device=torch.device("cpu"))
    exec("".join(nb["cells"][6]["source"]), ns)
    return ns["GrossPitaevskiiPINN"]


def load_paper_notebook_class(name="Gross_Pitaevskii_1D_Harmonic.ipynb", cell=6):
    """`GrossPitaevskiiPINN` of Notebooks/Paper/Gross_Pitaevskii_1D_Harmonic.ipynb cell 6."""
    from scipy.special import hermite
    with open(os.path.join(REF_ROOT, "Notebooks", "Paper", name)) as fh:
        nb = json.load(fh)
    ns = dict(torch=torch, nn=nn, np=np, math=math, hermite=hermite, device=torch.device("cpu"))
    exec("".join(nb["cells"][cell]["source"]), ns)
    return ns["GrossPitaevskiiPINN"]


def load_learnable_k_notebook_class():
    """`PINN` of Notebooks/Helmholtz/Helmholtz_learnable_k_and_bc.ipynb cell 3."""
    from torch.utils.checkpoint import checkpoint
    with open(os.path.join(REF_ROOT, "Notebooks", "Helmholtz", "Helmholtz_learnable_k_and_bc.ipynb")) as fh:
        nb = json.load(fh)
    ns = dict(torch=torch, nn=nn, np=np, math=math, checkpoint=checkpoint, device=torch.device("cpu"))
    exec("".join(nb["cells"][3]["source"]), ns)
    return ns["PINN"]


def load_pickle(relpath_under_refine):
    import pickle
    with open(os.path.join(REFINE, relpath_under_refine), "rb") as fh:
        return pickle.load(fh)
