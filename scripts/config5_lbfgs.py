"""BASELINE config 5 through the module API: 2-D GPE, optical-lattice + harmonic potential, [2,128x4,1],
torch.optim.LBFGS with the reference's settings (S/gross_pitaevskii_2D.py:670: max_iter 500, history 100,
tolerance_grad 1e-5, tolerance_change 1e-9) driving a full-batch closure over 8M collocation points per GPU
(64M on 8 GPUs).  Every closure evaluation = jet forward + moments (all-reduce #1) + backward (all-reduce #2), so all
ranks take the same line-search decisions.  Not a bench line: reports closure evaluations/s and points/s
(device time, max over ranks) for DESIGN.md / profiles/.

    python scripts/config5_lbfgs.py [--points 8388608] [--evals 12]
    torchrun --nproc-per-node N --master-addr 127.0.0.1 scripts/config5_lbfgs.py
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=8 * 1024 * 1024, help="collocation points per GPU")
    ap.add_argument("--evals", type=int, default=12, help="closure evaluations in the timed L-BFGS step (max_eval)")
    args = ap.parse_args()
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)                      # NCCL's banner must not land on stdout
    import gpe_b200
    from gpe_b200 import collocation, distributed as gdist

    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(1234)
    model = gpe_b200.GPE2D([2, 128, 128, 128, 128, 1], g=100.0, potential="lattice", semantics="pointwise")
    for m in model.modules():
        if isinstance(m, torch.nn.Linear):
            torch.nn.init.xavier_uniform_(m.weight)
            m.bias.data.fill_(0.01)
    model = model.to(dev)
    params = list(model.parameters())
    n = args.points
    X = collocation.latin_hypercube(n, 2, -6.0, 6.0, seed=1234 + rank, device=dev).requires_grad_(True)
    Xb = collocation.square_boundary(500, -6.0, 6.0, seed=1234, device=dev)
    losses = []

    def make_opt(max_eval):
        return torch.optim.LBFGS(params, max_iter=500, max_eval=max_eval, tolerance_grad=1e-5, tolerance_change=1e-9,
                                 history_size=100)

    def closure():
        for p in params:
            p.grad = None
        loss = model.loss(X, Xb, None)
        loss.backward()
        gdist.all_reduce_gradients(params)
        losses.append(loss.detach())
        return loss

    def sync():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    make_opt(3).step(closure)          # warm-up: allocator, plans, NCCL
    sync()
    n_warm = len(losses)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    make_opt(args.evals).step(closure)
    b.record()
    sync()
    evals = len(losses) - n_warm
    ms = torch.tensor([a.elapsed_time(b)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms)
    traj = [float(v) for v in losses]
    same = True
    if world > 1:                       # every rank must have seen the same loss sequence (same decisions)
        t = torch.tensor(traj, device=dev, dtype=torch.float64)
        lo, hi = t.clone(), t.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        same = bool(torch.equal(lo, hi))
    if rank == 0:
        line = {"workload": "BASELINE config 5: 2-D GPE, V = cos^2(kx) + cos^2(ky) + (x^2+y^2)/2, eta=100, [2,128x4,1], "
                            "L-BFGS (max_iter 500, history 100, tol_grad 1e-5) full-batch closure",
                "n_gpus": world, "points_per_gpu": n, "points_total": n * world, "closure_evals": evals,
                "ms_per_closure": ms / evals, "closure_evals_per_s": evals / ms * 1e3,
                "points_per_s": n * world * evals / ms * 1e3, "collectives_per_closure": 2 if world > 1 else 0,
                "loss_first": traj[n_warm], "loss_last": traj[-1], "loss_decreased": traj[-1] < traj[0],
                "all_ranks_same_losses": same}
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
