"""Epoch time of the paper's curriculum stage (4000 points, [1,64,64,64,1], R/harmonic_pinn_simulation.py main) through
gpe_b200.training.train_gpe_model: eager epochs vs one CUDA-graph replay per epoch.  Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
from gpe_b200 import training  # noqa: E402

lb, ub, n, layers, epochs = -10.0, 10.0, 4000, [1, 64, 64, 64, 1], 600
X = np.linspace(lb, ub, n).reshape(-1, 1)
torch.manual_seed(0)
from gpe_b200 import FinalGPE1D  # noqa: E402
init = FinalGPE1D(layers, mode=0, gamma=0.0)
init.apply(lambda m: training.advanced_initialization(m, 0))
state = {k: v.clone() for k, v in init.state_dict().items()}
out = {"workload": "one gamma stage, 4000 points, [1,64,64,64,1], %d epochs, lr 1e-5" % epochs}
for name, graph in (("eager", False), ("graph", True)):
    for rep in range(2):            # first repetition warms allocator / module load
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = training.train_gpe_model([0.0], [0], 3, X, lb, ub, layers, epochs, 1e-12, 0.01, lr=1e-5, verbose=False,
                                       init_state=state, use_graph=graph, check_every=100)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    out[name + "_ms_per_epoch"] = dt / epochs * 1e3
    out[name + "_mu"] = res[1][0][0][1]
out["speedup"] = out["eager_ms_per_epoch"] / out["graph_ms_per_epoch"]
print(json.dumps(out))
