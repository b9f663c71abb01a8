"""Sharded == single-GPU over NCCL: every rank evaluates the config-4 loss and gradient
once on the WHOLE collocation set by itself (a one-rank process group switches the collectives off) and once on its
contiguous shard with the two all-reduces (moments, packed gradient); loss, mu and the gradient must agree within the
fp32 reduction-order tolerance, and all ranks must hold bit-identical reduced gradients.

    torchrun --nproc-per-node N --master-addr 127.0.0.1 scripts/nccl_equivalence.py [--points 1000003]
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
    ap.add_argument("--points", type=int, default=1000003, help="total collocation points (ragged on purpose)")
    args = ap.parse_args()
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    import gpe_b200
    from gpe_b200 import collocation, distributed as gdist

    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    solo = None
    for r in range(world):                      # every rank takes part in the creation of every group
        g = dist.new_group([r])
        if r == rank:
            solo = g
    torch.manual_seed(1234)
    model = gpe_b200.GPE2D([2, 128, 128, 128, 128, 1], g=100.0, potential="harmonic", semantics="pointwise")
    for m in model.modules():
        if isinstance(m, torch.nn.Linear):
            torch.nn.init.xavier_uniform_(m.weight)
            m.bias.data.fill_(0.01)
    model = model.to(dev)
    params = list(model.parameters())
    X_all = collocation.latin_hypercube(args.points, 2, -6.0, 6.0, seed=1234, device=dev)   # same seed: same set on every rank
    Xb = collocation.square_boundary(500, -6.0, 6.0, seed=1234, device=dev)

    def evaluate(X, group):
        model.group = group
        for p in params:
            p.grad = None
        X = X.clone().requires_grad_(True)
        loss = model.loss(X, Xb, None)
        _, _, mu = model.pde_loss(X, None)
        loss.backward()
        gdist.all_reduce_gradients(params, group)
        return float(loss), float(mu), torch.cat([p.grad.reshape(-1) for p in params]).double()

    def boundary_only(group):
        """Gradient of the replicated boundary term alone (it must be counted once, not `world` times --
        in the full gradient the Riesz term hides it)."""
        model.group = group
        for p in params:
            p.grad = None
        model.boundary_loss(Xb, None).backward()
        gdist.all_reduce_gradients(params, group)
        return torch.cat([p.grad.reshape(-1) for p in params]).double()

    loss1, mu1, g1 = evaluate(X_all, solo)
    b1 = boundary_only(solo)
    lo, hi = gdist.shard_bounds(args.points, rank, world)
    lossN, muN, gN = evaluate(X_all[lo:hi].contiguous(), None)
    bN = boundary_only(None)
    rel = float((gN - g1).norm() / g1.norm())
    rel_b = float((bN - b1).norm() / b1.norm())
    gmin, gmax = gN.clone(), gN.clone()
    dist.all_reduce(gmin, op=dist.ReduceOp.MIN)
    dist.all_reduce(gmax, op=dist.ReduceOp.MAX)
    worst = torch.tensor([abs(lossN - loss1) / abs(loss1), abs(muN - mu1) / abs(mu1), rel, rel_b], device=dev, dtype=torch.float64)
    dist.all_reduce(worst, op=dist.ReduceOp.MAX)
    ok = bool(worst[0] < 1e-5 and worst[1] < 1e-6 and worst[2] < 1e-5 and worst[3] < 1e-6 and torch.equal(gmin, gmax))
    if rank == 0:
        line = {"n_gpus": world, "points": args.points, "shard_sizes": [gdist.shard_bounds(args.points, r, world)[1] -
                                                                       gdist.shard_bounds(args.points, r, world)[0] for r in range(world)],
                "loss_single": loss1, "loss_sharded": lossN, "mu_single": mu1, "mu_sharded": muN,
                "rel_loss": float(worst[0]), "rel_mu": float(worst[1]), "rel_grad": float(worst[2]),
                "rel_grad_boundary_term_only": float(worst[3]),
                "ranks_hold_identical_gradients": bool(torch.equal(gmin, gmax)), "ok": ok}
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
