"""Benchmark of the fused B200 PINN train step (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config 1..5]

Default workload (config.workload) = BASELINE config 4 -- 2-D GPE, harmonic trap, eta=100, tanh MLP
[2,128,128,128,128,1], N_f = 16M Latin-hypercube collocation points PER GPU (weak scaling: ranks own
independent blocks of points, parameters replicated) + 500 boundary points, loss = 10*bc + pde(+L_f
+L_lambda) + Riesz (S/gross_pitaevskii_2D.py:215-242, pointwise semantics).  One step = one full
evaluation of the loss and its parameter gradient through the public module API
(`model.loss(...)`, `.backward()`, gradient all-reduce for N>1).  `--config 1|2|3|5` selects the other
BASELINE shapes (parity-test cases whose lines are committed under profiles/; the driver runs the default).

  value     points/s, whole job, inputs resident in HBM, optimizer excluded
  strong    (config 4, same run) the 16M-point set of BASELINE config 4 SHARDED over the N ranks
            (gdist.shard_bounds): ms per step, points/s and efficiency against this run's own
            16M-points-on-one-GPU step time
  e2e       same metric through the same API from HOST buffers: per step the collocation block is copied
            from pinned host memory, the optimizer steps, and the loss is read back
  roofline  the slowest tcgen05 kernel of the step (stash forward / reverse sweep / dW contraction / fused
            backward), timed live with CUDA events on the launch stream (gpe_tc_profile); bound "tensor";
            achieved = algorithmic FLOPs (2 W^2 S (L-1) per point per contraction, S = 1+2d)
            / that time; peak = measured bf16 dense TFLOP/s of MEASURED_PEAKS.json (TF32 runs at half that rate
            and the 3xTF32 split issues 3 MMAs per product: both stated); traffic = ncu DRAM bytes of one launch
            from profiles/r2_ncu_traffic.json (written by scripts/ncu_digest.py with the git sha it profiled)
  gpu_launches  kernels launched in the timed region, COUNTED with torch.profiler (CUPTI) on one extra step
  cpu_baseline  the oracle port of the reference classes on the host cores, bounded sample, all threads and 1 thread

`--impl reference` times that CPU port alone (the reference is pure PyTorch; /root/reference does not
exist on the GPU box, so the restated classes in oracle/ stand in -- kind "port").
"""
import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

METRIC = "GPE PINN train-step collocation pts/sec (fwd+Laplacian+bwd)"
UNIT = "points/s"
N_BC = 500


def flops_per_point(layers, sym=False, streams=None):
    """F_fwd = 2 d W (1 + sym) + 2 W^2 S (L-1) + 2 W S with S = 1 + 2d (+1 mirrored value stream);
    F_step = 3 F_fwd."""
    d, W, L = layers[0], layers[1], len(layers) - 2
    S = ((1 + 2 * d) + (1 if sym else 0)) if streams is None else streams
    f_fwd = 2 * d * W * (2 if sym else 1) + 2 * W * W * S * (L - 1) + 2 * W * S
    return f_fwd, 3 * f_fwd


def init_weights(model):
    import torch.nn as nn
    torch.manual_seed(1234)
    for m in model.modules():
        if isinstance(m, nn.Linear):           # S/gross_pitaevskii_2D.py:614-618
            nn.init.xavier_uniform_(m.weight)
            m.bias.data.fill_(0.01)
    return model


def _perimeter(lo, hi, per_edge, seed):
    gen = torch.Generator().manual_seed(seed)
    t = lo + (hi - lo) * torch.rand(4, per_edge, generator=gen)
    lo_c, hi_c = torch.full_like(t[0], lo), torch.full_like(t[0], hi)
    return torch.cat([torch.stack([t[0], lo_c], 1), torch.stack([t[1], hi_c], 1), torch.stack([lo_c, t[2]], 1), torch.stack([hi_c, t[3]], 1)])


# --------------------------------------------------------------------------------------------------
# workloads: the five BASELINE configs.  `make(pkg, n, dev, seed, gpu)` returns
# (model, loss_fn(X) -> scalar loss, X) with pkg = gpe_b200 (CUDA) or oracle.pinn_oracle (CPU port)
# --------------------------------------------------------------------------------------------------
class Workload:
    TABLE = {
        1: dict(layers=[1, 100, 100, 100, 1], n=10_000, cpu_n=10_000, sym=True, domain=(-10.0, 10.0), opt="adam", lr=1e-3, n_bc=2,
                text="BASELINE config 1: 1-D GPE ground state, harmonic trap, gamma=1, p=3, tanh MLP [1,100,100,100,1], N_f=10k LHS "
                     "points, loss = pde + 10 bc + 20 norm + 5 sym (Gross_Pitaevskii_1D_power_Test.ipynb cell 10:88-103), Adam + clip 1.0"),
        2: dict(layers=[1, 128, 128, 128, 128, 1], n=1_000_000, cpu_n=100_000, sym=True, domain=(-10.0, 10.0), opt="adam", lr=1e-3, n_bc=2,
                text="BASELINE config 2: 1-D GPE eta=50, first excited mode (symmetry loss), tanh MLP [1,128x4,1], N_f=1M LHS points, "
                     "five-term total_loss (S/gross_pitaevskii_1D_Different_Optimizers.py:377-440), Adam"),
        3: dict(layers=[2, 100, 100, 100, 1], n=4_000_000, cpu_n=100_000, sym=False, domain=(0.0, math.pi), opt="adam", lr=1e-2, n_bc=400,
                text="BASELINE config 3: 2-D Helmholtz sin(pi x) sin(pi y), k = sqrt(2) pi, tanh MLP [2,100x3,1], N_f=4M LHS points + 400 "
                     "boundary points, compute_loss = pde + data (S/helmholtz_2D.py:110-132), pointwise semantics, Adam lr 0.01"),
        4: dict(layers=[2, 128, 128, 128, 128, 1], n=16 * 1024 * 1024, cpu_n=50_000, sym=False, domain=(-6.0, 6.0), opt="adam", lr=1e-3, n_bc=N_BC,
                text="BASELINE config 4: 2-D GPE harmonic trap eta=100, tanh MLP [2,128x4,1], N_f=16M LHS per GPU, "
                     "loss = 10*bc + pde(+L_f+L_lambda) + Riesz, pointwise semantics"),
        5: dict(layers=[2, 128, 128, 128, 128, 1], n=8 * 1024 * 1024, cpu_n=50_000, sym=False, domain=(-6.0, 6.0), opt="lbfgs", lr=1.0, n_bc=N_BC,
                text="BASELINE config 5: 2-D GPE, V = cos^2(kx) + cos^2(ky) + (x^2+y^2)/2, eta=100, tanh MLP [2,128x4,1], N_f=8M LHS per GPU "
                     "(64M on 8), L-BFGS full-batch closure (max_iter 500, history 100, tol_grad 1e-5: S/gross_pitaevskii_2D.py:670); "
                     "one step = one closure evaluation"),
    }

    def __init__(self, cfg):
        self.cfg = cfg
        self.__dict__.update(self.TABLE[cfg])
        self.d = self.layers[0]

    def points(self, n, dev, seed):
        """Latin-hypercube collocation block [n,d] on `dev` (generated where it lives; CPU for the port)."""
        if dev.type == "cuda":
            from gpe_b200 import collocation
            return collocation.latin_hypercube(n, self.d, self.domain[0], self.domain[1], seed=seed, device=dev)
        gen = torch.Generator().manual_seed(seed)
        cols = [(torch.randperm(n, generator=gen).double() + torch.rand(n, generator=gen, dtype=torch.float64)) / n for _ in range(self.d)]
        return (self.domain[0] + (self.domain[1] - self.domain[0]) * torch.stack(cols, 1)).float()

    def make(self, pkg, n, dev, seed, gpu):
        lo, hi = self.domain
        X = self.points(n, dev, seed).requires_grad_(True)
        if self.cfg == 1:
            model = init_weights(pkg.RootGPE1D(self.layers, mode=0, gamma=1.0)).to(dev)
            bp = torch.tensor([[lo], [hi]], device=dev)
            bv = torch.zeros(2, 1, device=dev)
            dx = (hi - lo) / n

            def loss_fn(Xc):
                u_pred = model.forward(Xc)
                pde, _, _, full_u = model.pde_loss(3, Xc, u_pred, 1.0, "harmonic")
                return (pde + 10.0 * model.boundary_loss(bp, bv) + 20.0 * model.normalization_loss(full_u, dx)
                        + 5.0 * model.symmetry_loss(Xc, lo, hi))
        elif self.cfg == 2:
            model = init_weights(pkg.OptGPE1D(self.layers, g=50.0, mode=1)).to(dev)
            bp = torch.tensor([[lo], [hi]], device=dev)
            bv = torch.zeros(2, 1, device=dev)

            def loss_fn(Xc):
                return model.total_loss(Xc, bp, bv, 50.0, lo, hi, [1.0] * 5, "harmonic")[0]
        elif self.cfg == 3:
            kw = dict(semantics="pointwise") if gpu else {}
            model = init_weights(pkg.Helmholtz2D(self.layers, activation="tanh", **kw)).to(dev)
            Xu = _perimeter(lo, hi, 100, 7).to(dev)
            Uu = torch.zeros(400, 1, device=dev)
            k = math.sqrt(2.0) * math.pi
            if gpu:
                def loss_fn(Xc):
                    return model.compute_loss(Xc, Xu, Uu, k)
            else:
                from oracle import pinn_oracle as po

                def loss_fn(Xc):      # the port's class keeps the reference's [N,N] broadcast; the pointwise form is timed
                    u = model.forward(Xc)
                    gr = po.d_dx(u, Xc)
                    lap = po.d_dx(gr[:, [0]], Xc)[:, [0]] + po.d_dx(gr[:, [1]], Xc)[:, [1]]
                    q = k ** 2 * torch.sin(math.pi * Xc[:, [0]]) * torch.sin(math.pi * Xc[:, [1]])
                    return torch.mean((lap + k ** 2 * u - q) ** 2) + model.loss_data(Xu, Uu)
        else:
            pot = "harmonic" if self.cfg == 4 else "lattice"
            kw = dict(semantics="pointwise") if gpu else {}
            model = init_weights(pkg.GPE2D(self.layers, g=100.0, potential=pot, **kw)).to(dev)
            Xb = _perimeter(lo, hi, N_BC // 4, 1241).to(dev)
            if gpu:
                def loss_fn(Xc):
                    return model.loss(Xc, Xb, None)
            else:
                from oracle import pinn_oracle as po

                def loss_fn(Xc):
                    return po.gpe2d_pointwise_terms(model, Xc, Xb)[0]
        return model, loss_fn, X


# --------------------------------------------------------------------------------------------------
# CPU port of the reference (oracle) -- the reported baseline and the `--impl reference` arm
# --------------------------------------------------------------------------------------------------
def cpu_reference_steps(wl, n_points, steps, warmup, threads):
    from oracle import pinn_oracle as po
    torch.set_num_threads(threads)
    model, loss_fn, X = wl.make(po, n_points, torch.device("cpu"), 1234, gpu=False)
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        model.zero_grad(set_to_none=True)
        loss_fn(X).backward()
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    return times


def cpu_reference_semantics_steps(n_points, steps, threads):
    """The reference's own [N,N] semantics (S/gross_pitaevskii_2D.py:183-213 keeps `[N]` x `[N,1]` shapes): the restated
    class's `loss`, N <= 2000 so the matrix fits."""
    from oracle import pinn_oracle as po
    torch.set_num_threads(threads)
    model = init_weights(po.GPE2D([2, 128, 128, 128, 128, 1], g=100.0, potential="harmonic"))
    gen = torch.Generator().manual_seed(3)
    X = (-6.0 + 12.0 * torch.rand(n_points, 2, generator=gen)).requires_grad_(True)
    Xb = -6.0 + 12.0 * torch.rand(N_BC, 2, generator=gen)
    times = []
    for it in range(1 + steps):
        t0 = time.perf_counter()
        model.zero_grad(set_to_none=True)
        model.loss(X, Xb, torch.zeros(N_BC, 1)).backward()
        if it >= 1:
            times.append(time.perf_counter() - t0)
    return times


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = Workload(args.config)
    threads = os.cpu_count() or 1
    n = args.cpu_points or wl.cpu_n
    times = cpu_reference_steps(wl, n, args.steps, args.warmup, threads)
    ms = 1e3 * sum(times) / len(times)
    value = n / (ms / 1e3)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl.text, "baseline_config": wl.cfg, "layers": wl.layers, "points_per_step": n, "device": "cpu"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"{n} collocation points per step, oracle/pinn_oracle (restated reference class), "
                                       "fwd + nested autograd.grad + losses + backward"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
# clocks sampler (B200_PROFILING.md recipe)
# --------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.rows.append([c.strip() for c in ln.split(",")])

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 7:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def count_kernels(step):
    """Device activities one step launches, counted from the CUPTI records torch.profiler collects: (kernels of this
    repo, all kernels + memsets + memcpys)."""
    from torch.profiler import ProfilerActivity, profile
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        step()
        torch.cuda.synchronize()
    ours, total = 0, 0
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            total += 1
            if "gpe::" in ev.name or "tcp::" in ev.name:
                ours += 1
    return ours, total


# --------------------------------------------------------------------------------------------------
# the B200 arm
# --------------------------------------------------------------------------------------------------
def run_b200(args):
    # stdout carries exactly ONE JSON line: native libraries write to fd 1 behind python's back (NCCL prints its
    # "NCCL version ..." banner there at communicator creation), so fd 1 is pointed at stderr for the run and the line
    # goes to the saved descriptor
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    import torch.distributed as dist
    import gpe_b200
    from gpe_b200 import distributed as gdist, ops

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- gpe_b200 has no CPU path (use --impl reference for the CPU port)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    wl = Workload(args.config)
    n = args.points or wl.n
    model, loss_fn, X = wl.make(gpe_b200, n, dev, 1234 + rank, gpu=True)
    params = list(model.parameters())
    steps, warmup = args.steps, max(args.warmup, 3)

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    from gpe_b200.models import invalidate_caches

    def make_step(Xc):
        def step(events=None):
            for p in params:
                p.grad = None
            if events:
                events[0].record()
            invalidate_caches()      # a training step follows an optimizer step: nothing of the previous step's jets may be reused
            loss = loss_fn(Xc)
            if events:
                events[1].record()
            loss.backward()
            if events:
                events[2].record()
            gdist.all_reduce_gradients(params)
            return loss
        return step

    def timed(step, k):
        """ms per step of `k` steps (device time, max over ranks) + per-phase means of this rank."""
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(k)]
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sync_all()
        t0.record()
        for i in range(k):
            loss = step(ev[i])
        t1.record()
        sync_all()
        ms = torch.tensor([t0.elapsed_time(t1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        # the loss goes back as a float: a live loss tensor keeps its autograd graph (and AccumulateGrad nodes bound to this
        # stream) alive, which would invalidate a later CUDA-graph capture of the same parameters
        return (float(ms) / k, sum(e[0].elapsed_time(e[1]) for e in ev) / k, sum(e[1].elapsed_time(e[2]) for e in ev) / k, float(loss.detach()))

    step = make_step(X)

    # ---- FFMA peak (the fp32 roofline denominator is not in MEASURED_PEAKS.json: measure it here)
    peak = {}
    for mode, key in ((0, "ffma"), (1, "ffma2")):
        ops.ffma_probe(mode, 2000, dev)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(dev)
        a.record()
        fl = ops.ffma_probe(mode, 20000, dev)
        b.record()
        torch.cuda.synchronize(dev)
        peak[key] = fl / (a.elapsed_time(b) * 1e-3) / 1e12

    for _ in range(warmup):
        step()
    sync_all()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms_per_step, fwd_ms, bwd_ms, loss = timed(step, steps)
    clocks = sampler.stop() if rank == 0 else None
    value = world * n / (ms_per_step * 1e-3)

    # ---- small configurations are launch-bound: the same step (+ clip + Adam) as one CUDA-graph replay
    graph = None
    if n <= 200_000 and wl.opt == "adam" and world == 1:
        import gc
        from gpe_b200.graphs import GraphedStep
        for p in params:
            p.grad = None
        gc.collect()
        saved = {k: v.clone() for k, v in model.state_dict().items()}
        gopt = torch.optim.Adam(params, lr=wl.lr, capturable=True)

        def body():
            gopt.zero_grad(set_to_none=True)
            lossg = loss_fn(X)
            lossg.backward()
            torch.nn.utils.clip_grad_norm_(params, max_norm=1.0)
            gopt.step()
            return lossg
        gs = GraphedStep(body, gopt)
        for _ in range(5):
            gs()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(dev)
        a.record()
        reps = max(50, steps)
        for _ in range(reps):
            gs()
        b.record()
        torch.cuda.synchronize(dev)
        gms = a.elapsed_time(b) / reps
        graph = {"ms_per_step": gms, "value": n / (gms * 1e-3), "unit": UNIT,
                 "includes": "loss, backward, clip_grad_norm_, Adam (capturable) in one CUDA-graph replay", "replays_timed": reps}
        model.load_state_dict(saved)
        for p in params:
            p.grad = None

    # ---- strong scaling of BASELINE config 4 as it is written: ONE 16M-point set sharded over the ranks
    strong = None
    if wl.cfg == 4 and not args.no_strong:
        total = args.strong_points
        if world == 1:
            if n == total:
                strong = {"points_total": n, "points_per_gpu": n, "ms_per_step": ms_per_step, "value": value, "efficiency_vs_n1": 1.0,
                          "note": "one rank: the sharded set is the whole set"}
        else:
            lo, hi = gdist.shard_bounds(total, rank, world)
            Xs = wl.points(hi - lo, dev, 4321 + rank).requires_grad_(True)   # this rank's shard of the 16M-point set
            sstep = make_step(Xs)
            for _ in range(warmup):
                sstep()
            sms, sf, sb, _ = timed(sstep, steps)
            strong = {"points_total": total, "points_per_gpu": hi - lo, "ms_per_step": sms, "value": total / (sms * 1e-3),
                      "phase_ms": {"forward_and_losses": sf, "backward": sb},
                      # one GPU's time for the same 16M points = this run's own weak step (n points per rank) scaled to 16M
                      "efficiency_vs_n1": (ms_per_step * (total / n) / world) / sms,
                      "reference_n1_ms": ms_per_step * (total / n), "collectives_per_step": 2}
            del Xs, sstep

    # ---- end to end: host buffers in, loss out, optimizer included
    if wl.opt == "lbfgs":
        opt = torch.optim.LBFGS(params, max_iter=500, max_eval=steps + 2, tolerance_grad=1e-5, tolerance_change=1e-9, history_size=100)
    else:
        opt = torch.optim.Adam(params, lr=wl.lr)
    # Every step's collocation block comes from pinned host memory.  Two device buffers: the copy of step k+1 runs on a
    # copy stream while step k computes (what a loader feeding fresh points per step does); step k+1 waits for its copy,
    # and a buffer is not overwritten before the step that read it has finished.
    X_host = X.detach().cpu().pin_memory()
    X_bufs = [torch.empty_like(X_host, device=dev).requires_grad_(True) for _ in range(2)]
    copy_stream = torch.cuda.Stream(device=dev)
    copied = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    e2e_k = [0]

    def upload(i):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[i])
            with torch.no_grad():
                X_bufs[i].copy_(X_host, non_blocking=True)
            copied[i].record(copy_stream)

    for i in range(2):
        consumed[i].record()
    upload(0)

    def e2e_eval():
        cur = e2e_k[0] % 2
        e2e_k[0] += 1
        main = torch.cuda.current_stream(dev)
        main.wait_event(copied[cur])
        upload(1 - cur)                    # next evaluation's inputs, overlapping this one's kernels
        opt.zero_grad(set_to_none=True)
        loss_e = loss_fn(X_bufs[cur])
        loss_e.backward()
        consumed[cur].record(main)
        gdist.all_reduce_gradients(params)
        return loss_e

    saved = {k: v.clone() for k, v in model.state_dict().items()}
    evals = [0]
    if wl.opt == "lbfgs":
        def closure():
            evals[0] += 1
            return e2e_eval()
        torch.optim.LBFGS(params, max_iter=500, max_eval=3, tolerance_grad=1e-5, tolerance_change=1e-9, history_size=100).step(closure)
        sync_all()
        evals[0] = 0
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        last = float(opt.step(closure))     # the line search reads every closure value on the host
        torch.cuda.current_stream(dev).wait_event(copied[e2e_k[0] % 2])
        b.record()
        sync_all()
        n_evals = max(evals[0], 1)
    else:
        def e2e_step():
            loss_e = e2e_eval()
            if wl.cfg == 1:
                torch.nn.utils.clip_grad_norm_(params, max_norm=1.0)
            opt.step()
            return float(loss_e.detach())   # device -> host read of the step's result
        for _ in range(2):
            e2e_step()
        sync_all()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            last = e2e_step()
        torch.cuda.current_stream(dev).wait_event(copied[e2e_k[0] % 2])   # the K-th upload issued in the region ends inside it
        b.record()
        sync_all()
        n_evals = steps
    e2e_ms = torch.tensor([a.elapsed_time(b)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_value = world * n / (float(e2e_ms) / n_evals * 1e-3)
    model.load_state_dict(saved)

    # ---- per-kernel times of the tensor path (one extra, untimed step with event brackets) and the launch count
    tensor = os.environ.get("GPE_TENSOR", "1") != "0"
    tc_ms = None
    if tensor:
        ops.tc_profile(True)
        step()
        tc_ms = ops.tc_timings()
        ops.tc_profile(False)
        sync_all()
    ours, total_k = count_kernels(step) if rank == 0 or world > 1 else (0, 0)
    sync_all()

    if rank == 0:
        layers = wl.layers
        d, W, L = layers[0], layers[1], len(layers) - 2
        f_fwd, f_step = flops_per_point(layers, wl.sym)
        peaks_file = os.path.join(ROOT, "MEASURED_PEAKS.json")
        measured = json.load(open(peaks_file)) if os.path.exists(peaks_file) else {}
        fp32_peak = max(peak.values())
        tc_on = bool(tc_ms) and max(tc_ms.values()) > 0
        resident = tc_on and tc_ms.get("stash_forward", 1.0) == 0.0
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": wl.text, "baseline_config": wl.cfg, "layers": layers, "points_per_gpu": n,
                       "boundary_points": wl.n_bc, "parallelism": f"points-dp{world}",
                       "stash": "resident (forward keeps 2 KB per point and hidden matrix of pre-activation jets, backward does not recompute)"
                                if resident else "recomputed per chunk in the backward",
                       "l2_policy": ("inputs (points + jets/adjoints + 2 KB/point/matrix of stash per step) exceed the 126 MB L2"
                                     if n * (6 * d + 2048 * (L - 1)) > 126e6 else
                                     "the working set fits L2 and every step rewrites it: a small-N line measures launch latency, not bandwidth"),
                       "collectives_per_step": 0 if world == 1 else 2},
            "phase_ms": {"forward_and_losses": fwd_ms, "backward": bwd_ms},
            "step_tflops_algorithmic": f_step * n / (ms_per_step * 1e-3) / 1e12,
            # the whole step against the FP32 FFMA roofline the reference's arithmetic (fp32) would be bound by: > 1 because
            # the contractions run as 3xTF32 on the tensor cores
            "step_frac_of_fp32_ffma_peak": f_step * n / (ms_per_step * 1e-3) / 1e12 / fp32_peak,
            "roofline": None,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": X_host.numel() * 4 * world, "d2h_bytes_per_step": 4 * world,
                    "includes": "pinned-host -> device copy of every step's collocation block (double-buffered on a copy stream, "
                                "overlapping the previous step), loss, backward, all-reduce, optimizer step"
                                + (" (clip_grad_norm_ 1.0)" if wl.cfg == 1 else "") + ", loss read-back"
                                + (f"; L-BFGS: {n_evals} closure evaluations in the timed step" if wl.opt == "lbfgs" else ""),
                    "last_loss": last},
            "gpu_launches": ours * steps,
            "gpu_launches_per_step": {"gpe_kernels": ours, "all_device_activities": total_k, "how": "torch.profiler (CUPTI) records of one step"},
            "clocks": clocks, "loss": loss,
        }
        if graph:
            line["graph_replay"] = graph
        if strong:
            line["strong"] = strong
            line["config"]["headline"] = ("`value` = weak scaling (16M points per GPU, round-over-round continuity); `strong` = BASELINE "
                                          "config 4 as written (one 16M-point set sharded over the ranks)")
        f_contr = 2 * W * W * (1 + 2 * d) * (L - 1)          # one hidden-layer contraction sweep per point (algorithmic)
        bf16 = measured.get("bf16_tflops_sustained") or 1400.0
        if tensor and tc_on:
            kname = max(tc_ms, key=tc_ms.get)
            kms = tc_ms[kname]
            alg = {"stash_forward": f_contr, "reverse_sweep": f_contr, "dw_contraction": f_contr, "fused_backward": 2 * f_contr}[kname]
            sym = {"stash_forward": f"tcp::jet_fwd2_tc_kernel<{d},true,ACT_TANH>", "reverse_sweep": (f"tcp::jet_rev_tc_kernel<{d},ACT_TANH>" if os.environ.get("GPE_TC_TWO_TILES_REV", "1")[:1] == "0"
                                     else f"tcp::jet_rev2_tc_kernel<{d},ACT_TANH>"),
                   "dw_contraction": f"tcp::dw_tc_kernel<{d},ACT_TANH>", "fused_backward": f"tcp::bwd_fused_tc_kernel<{d},ACT_TANH>"}[kname]
            achieved = alg * n / (kms * 1e-3) / 1e12
            # image blocks of 2 KB per point and hidden matrix l = 1..L-1: st_l written by the stash forward, read by the
            # reverse sweep (all) and the dW contraction (l <= L-2); zbar_l written by the reverse sweep, read by dW
            hbm_bytes = {"stash_forward": 4 * d + 2048 * (L - 1), "reverse_sweep": 4 * d + 4 * (2 + d) + 2 * 2048 * (L - 1),
                         "dw_contraction": 4 * d + 2048 * (L - 1) + 2048 * (L - 2),
                         "fused_backward": 4 * d + 4 * (2 + d) + 2048 * (L - 1) + 2048 * (L - 2)}[kname]
            chunk = int(os.environ.get("GPE_TC_CHUNK", 1048576))
            launches = 1 if kname in ("stash_forward", "fused_backward") else (n + chunk - 1) // chunk
            tfile = os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")
            tinfo = json.load(open(tfile)) if os.path.exists(tfile) else {}
            trow = tinfo.get("kernels", {}).get(kname)
            line["roofline"] = {
                "bound": "tensor", "kernel": sym, "achieved": achieved, "peak": bf16, "unit": "TFLOP/s", "frac": achieved / bf16,
                "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained" if measured.get("bf16_tflops_sustained") else "fallback 1.4 PFLOP/s",
                "tf32_peak": bf16 / 2, "effective_3xtf32_peak": bf16 / 6, "frac_of_3xtf32_peak": achieved / (bf16 / 6),
                # what THIS kernel's MMA schedule could reach: per 32-feature chunk the sweeps issue 4 tf32 + 4 bf16 MMAs of equal
                # duration (8 slots: tf32 peak x 4/8), the dW contraction 12 tf32 MMAs (tf32 peak x 4/12)
                "mma_slots_per_chunk": 12 if kname == "dw_contraction" else 8,
                "effective_peak_of_this_schedule": (bf16 / 6) if kname == "dw_contraction" else (bf16 / 4),
                "frac_of_this_schedule": achieved / ((bf16 / 6) if kname == "dw_contraction" else (bf16 / 4)),
                "algorithmic_flops_per_point": alg, "executed_tensor_flops_per_point": 3 * 2 * W * W * (2 + d) * (L - 1) * (alg // f_contr),
                # same multiply-adds in every variant; the sweeps issue their two correction products as bf16 kind::f16 MMAs
                # (K = 16: 8 MMA issue slots per 32-feature chunk instead of 12), the dW contraction keeps three tf32 products
                "correction_products": {"stash_forward": "bf16 (kind::f16)", "reverse_sweep": "bf16 (kind::f16)", "dw_contraction": "tf32",
                                        "fused_backward": "bf16 (sweep) + tf32 (dW)"}[kname],
                "kernel_ms_per_step": kms, "kernel_launches_per_step": launches, "tensor_step_ms": tc_ms,
                "traffic": (trow or {}).get("dram_bytes_per_launch"),
                "traffic_source": ({"file": "profiles/r2_ncu_traffic.json", "git_sha": tinfo.get("git_sha"),
                                    "points_per_launch": trow.get("points_per_launch"), "dram_bytes_per_point": trow.get("dram_bytes_per_point")}
                                   if trow else None),
                "algorithmic_hbm_bytes_per_point": hbm_bytes, "hbm_gbs_achieved": hbm_bytes * n / (kms * 1e-3) / 1e9,
                "hbm_gbs_measured": measured.get("hbm_gbs"), "fp32_ffma_peak_tflops": fp32_peak}
        else:
            bwd_alg = 2 * f_fwd * n / (bwd_ms * 1e-3) / 1e12
            line["roofline"] = {"bound": "fp32", "kernel": f"jet_bwd_kernel<{d},...>", "achieved": bwd_alg, "peak": fp32_peak,
                                "unit": "TFLOP/s", "frac": bwd_alg / fp32_peak,
                                "peak_source": "FFMA/FFMA2 probe kernel timed in this run (MEASURED_PEAKS.json has no fp32 figure); "
                                               f"ffma={peak['ffma']:.1f} ffma2={peak['ffma2']:.1f} TFLOP/s",
                                "algorithmic_flops_per_point": 2 * f_fwd, "traffic": None,
                                "tensor_peak_tflops_bf16_measured": measured.get("bf16_tflops_sustained"),
                                "hbm_gbs_measured": measured.get("hbm_gbs")}
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            cn = args.cpu_points or wl.cpu_n
            times = cpu_reference_steps(wl, cn, 3, 1, threads)
            v = cn / (sum(times) / len(times))
            n1 = max(cn // 8, 2000)
            t1 = cpu_reference_steps(wl, n1, 2, 1, 1)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": f"{cn} collocation points per step x 3 steps (1 warm-up), oracle/pinn_oracle "
                                              "(restated reference class) on the host cores",
                                    "one_thread": {"value": n1 / (sum(t1) / len(t1)), "cores": 1, "sample": f"{n1} points x 2 steps"}}
            if wl.cfg == 4:
                tr = cpu_reference_semantics_steps(1500, 2, threads)
                line["cpu_baseline"]["reference_semantics"] = {
                    "value": 1500 / (sum(tr) / len(tr)), "cores": threads,
                    "sample": "the reference's own [N,N] broadcast (restated class `loss`), N = 1500 x 2 steps"}
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=4, choices=[1, 2, 3, 4, 5], help="BASELINE config (default 4: the driver's line)")
    ap.add_argument("--points", type=int, default=0, help="collocation points per GPU (0: the config's own size)")
    ap.add_argument("--cpu-points", type=int, default=0, help="bounded sample for the CPU port (0: the config's default)")
    ap.add_argument("--strong-points", type=int, default=16 * 1024 * 1024, help="size of the sharded set of the `strong` record")
    ap.add_argument("--no-strong", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
