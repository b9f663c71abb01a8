#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
python - <<'PY'
import sys, torch, numpy as np
sys.path.insert(0, "gross-pitaevskii-eigenvalue-problem_b200"); sys.path.insert(0, ".")
import gpe_b200
from oracle import pinn_oracle as po
torch.manual_seed(0)
layers = [2, 200, 200, 200, 1]
ref = po.Helmholtz2D(layers).double()
gpu = gpe_b200.Helmholtz2D(layers, semantics="reference"); gpu.load_state_dict({k: v.float() for k, v in ref.state_dict().items()}); gpu = gpu.cuda()
X = torch.rand(500, 2, dtype=torch.float64) * np.pi; Xu = torch.rand(80, 2, dtype=torch.float64) * np.pi; U = torch.rand(80, 1, dtype=torch.float64)
k = float(np.sqrt(2) * np.pi)
lr = ref.compute_loss(X.clone().requires_grad_(True), Xu, U, k); lr.backward()
lg = gpu.compute_loss(X.float().cuda().requires_grad_(True), Xu.float().cuda(), U.float().cuda(), k); lg.backward()
gr = torch.cat([p.grad.reshape(-1) for p in ref.parameters()]); gg = torch.cat([p.grad.reshape(-1) for p in gpu.parameters()]).double().cpu()
print("helmholtz W=200 (reference default, leaky relu): loss rel", abs(float(lg) - float(lr)) / abs(float(lr)), "grad rel", float((gg - gr).norm() / gr.norm()))
PY
