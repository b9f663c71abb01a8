"""Benchmark of the fused B200 PINN train step (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Workload (config.workload): BASELINE config 4 -- 2-D GPE, harmonic trap, eta=100, tanh MLP
[2,128,128,128,128,1], N_f = 16M Latin-hypercube collocation points PER GPU (weak scaling: ranks own
independent blocks of points, parameters replicated) + 500 boundary points, loss = 10*bc + pde(+L_f
+L_lambda) + Riesz (S/gross_pitaevskii_2D.py:215-242, pointwise semantics).  One step = one full
evaluation of the loss and its parameter gradient through the public module API
(`model.loss(...)`, `.backward()`, gradient all-reduce for N>1).

  value     points/s, whole job, inputs resident in HBM, optimizer excluded
  e2e       same metric through the same API from HOST buffers: per step the collocation block is copied
            from pinned host memory, Adam steps, and the loss is read back
  roofline  tensor path (default): the largest of the three tcgen05 backward kernels (stash forward, reverse
            sweep, dW contraction), timed live with CUDA events on the launch stream (gpe_tc_profile);
            bound "tensor"; achieved = algorithmic FLOPs (2 W^2 S (L-1) per point per
            contraction, S = 1+2d) / that time; peak = measured bf16 dense TFLOP/s of MEASURED_PEAKS.json
            (TF32 runs at half that rate and the 3xTF32 split issues 3 MMAs per product: both stated).
            With GPE_TENSOR=0: jet_bwd_kernel, FP32-FFMA bound, as in the first round-1 line
  cpu_baseline  the oracle port of the reference classes on the host cores, bounded sample

`--impl reference` times that CPU port alone (the reference is pure PyTorch; /root/reference does not
exist on the GPU box, so the restated classes in oracle/ stand in -- kind "port").
"""
import argparse
import json
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

LAYERS = [2, 128, 128, 128, 128, 1]
ETA = 100.0
DOMAIN = (-6.0, 6.0)
N_BC = 500
METRIC = "GPE PINN train-step collocation pts/sec (fwd+Laplacian+bwd)"
UNIT = "points/s"


def flops_per_point(layers=LAYERS, streams=None):
    """F_fwd = 2 d W + 2 W^2 S (L-1) + 2 W S with S = 1 + 2d; F_step = 3 F_fwd."""
    d, W, L = layers[0], layers[1], len(layers) - 2
    S = (1 + 2 * d) if streams is None else streams
    f_fwd = 2 * d * W + 2 * W * W * S * (L - 1) + 2 * W * S
    return f_fwd, 3 * f_fwd


def init_weights(model):
    import torch.nn as nn
    torch.manual_seed(1234)
    for m in model.modules():
        if isinstance(m, nn.Linear):           # S/gross_pitaevskii_2D.py:614-618
            nn.init.xavier_uniform_(m.weight)
            m.bias.data.fill_(0.01)
    return model


# --------------------------------------------------------------------------------------------------
# CPU port of the reference (oracle) -- the reported baseline and the `--impl reference` arm
# --------------------------------------------------------------------------------------------------
def cpu_reference_steps(n_points, steps, warmup, threads):
    from oracle import pinn_oracle as po
    torch.set_num_threads(threads)
    model = init_weights(po.GPE2D(LAYERS, g=ETA, potential="harmonic"))
    gen = torch.Generator().manual_seed(1234)
    X = (DOMAIN[0] + (DOMAIN[1] - DOMAIN[0]) * torch.rand(n_points, 2, generator=gen)).requires_grad_(True)
    Xb = DOMAIN[0] + (DOMAIN[1] - DOMAIN[0]) * torch.rand(N_BC, 2, generator=gen)
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        model.zero_grad(set_to_none=True)
        total, *_ = po.gpe2d_pointwise_terms(model, X, Xb)
        total.backward()
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    return times


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n = args.cpu_points
    times = cpu_reference_steps(n, args.steps, args.warmup, threads)
    ms = 1e3 * sum(times) / len(times)
    value = n / (ms / 1e3)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "BASELINE config 4: 2-D GPE harmonic trap eta=100, tanh MLP [2,128x4,1], pointwise semantics",
                       "layers": LAYERS, "points_per_step": n, "device": "cpu"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"{n} of the 16M collocation points per step, oracle/pinn_oracle.GPE2D "
                                       "(restated reference class), fwd + 4x autograd.grad + losses + backward"},
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


# --------------------------------------------------------------------------------------------------
# the B200 arm
# --------------------------------------------------------------------------------------------------
KERNELS_PER_STEP = 12   # FFMA path: 2x(pack_weights + jet_fwd) + gram_moments + gram_adjoints + 2x(pack + jet_bwd + finish_grads)
TC_CHUNK = 262144       # points per (stash forward, reverse sweep, dW) round of the tensor-path backward


def kernels_per_step(n, tensor, resident=False):
    if not tensor:
        return KERNELS_PER_STEP
    chunks = (n + TC_CHUNK - 1) // TC_CHUNK
    if resident:   # the loss forward writes the stash: 2 kernels per chunk in the backward
        return 2 + 3 + 2 + 3 + 4 + 2 * chunks
    # boundary fwd (pack + jet_fwd) + main fwd (pack + pack_wimg + fwd_tc) + moments + adjoints
    # + boundary bwd (pack + jet_bwd + finish) + main bwd (pack + 2 pack_wimg + 3 per chunk + finish)
    # (torch's own elementwise/reduction launches for the fp64 scalar algebra are not counted)
    return 2 + 3 + 2 + 3 + 4 + 3 * chunks


def run_b200(args):
    # stdout carries exactly ONE JSON line: native libraries write to fd 1 behind python's back (NCCL prints its
    # "NCCL version ..." banner there at communicator creation), so fd 1 is pointed at stderr for the run and the line
    # goes to the saved descriptor
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    import torch.distributed as dist
    import gpe_b200
    from gpe_b200 import collocation, distributed as gdist, ops

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- gpe_b200 has no CPU path (use --impl reference for the CPU port)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = args.points
    model = init_weights(gpe_b200.GPE2D(LAYERS, g=ETA, potential="harmonic", semantics="pointwise")).to(dev)
    params = list(model.parameters())
    X = collocation.latin_hypercube(n, 2, DOMAIN[0], DOMAIN[1], seed=1234 + rank, device=dev).requires_grad_(True)
    Xb = collocation.square_boundary(N_BC, DOMAIN[0], DOMAIN[1], seed=1234, device=dev)

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def step(events=None):
        for p in params:
            p.grad = None
        if events:
            events[0].record()
        loss = model.loss(X, Xb, None)
        if events:
            events[1].record()
        loss.backward()
        if events:
            events[2].record()
        gdist.all_reduce_gradients(params)
        return loss

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

    for _ in range(max(args.warmup, 3)):
        step()
    sync_all()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    t_begin, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    t_begin.record()
    for k in range(args.steps):
        loss = step(ev[k])
    t_end.record()
    sync_all()
    clocks = sampler.stop() if rank == 0 else None
    total_ms = torch.tensor([t_begin.elapsed_time(t_end)], device=dev, dtype=torch.float64)
    fwd_ms = sum(e[0].elapsed_time(e[1]) for e in ev) / args.steps
    bwd_ms = sum(e[1].elapsed_time(e[2]) for e in ev) / args.steps
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    ms_per_step = float(total_ms) / args.steps
    value = world * n / (ms_per_step * 1e-3)

    # ---- end to end: host buffers in, loss out, Adam included
    opt = torch.optim.Adam(params, lr=1e-3)
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

    def e2e_step():
        cur = e2e_k[0] % 2
        e2e_k[0] += 1
        main = torch.cuda.current_stream(dev)
        main.wait_event(copied[cur])
        upload(1 - cur)                    # next step's inputs, overlapping this step's kernels
        opt.zero_grad(set_to_none=True)
        loss = model.loss(X_bufs[cur], Xb, None)
        loss.backward()
        consumed[cur].record(main)
        gdist.all_reduce_gradients(params)
        opt.step()
        return float(loss.detach())   # device -> host read of the step's result

    saved = {k: v.clone() for k, v in model.state_dict().items()}
    for _ in range(2):
        e2e_step()
    sync_all()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.steps):
        last = e2e_step()
    torch.cuda.current_stream(dev).wait_event(copied[e2e_k[0] % 2])   # the K-th upload issued in the region ends inside it
    b.record()
    sync_all()
    e2e_ms = torch.tensor([a.elapsed_time(b)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_value = world * n / (float(e2e_ms) / args.steps * 1e-3)
    model.load_state_dict(saved)

    # ---- per-kernel times of the tensor-path backward, one extra (untimed) step with event brackets
    tensor = os.environ.get("GPE_TENSOR", "1") != "0"
    tc_ms = None
    if tensor:
        ops.tc_profile(True)
        step()
        tc_ms = ops.tc_timings()
        ops.tc_profile(False)
        sync_all()

    if rank == 0:
        f_fwd, f_step = flops_per_point()
        f_fwd4, _ = flops_per_point(streams=4)
        bwd_alg = 2 * f_fwd * n / (bwd_ms * 1e-3) / 1e12
        peaks_file = os.path.join(ROOT, "MEASURED_PEAKS.json")
        measured = json.load(open(peaks_file)) if os.path.exists(peaks_file) else {}
        fp32_peak = max(peak.values())
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": "BASELINE config 4: 2-D GPE harmonic trap eta=100, tanh MLP [2,128x4,1], N_f=16M LHS per GPU, "
                                   "loss = 10*bc + pde(+L_f+L_lambda) + Riesz, pointwise semantics",
                       "layers": LAYERS, "points_per_gpu": n, "boundary_points": N_BC, "parallelism": f"points-dp{world}",
                       "stash": "resident (forward keeps 6 KB/point of pre-activation jets, backward does not recompute)"
                                if (tensor and tc_ms and tc_ms.get("stash_forward", 1.0) == 0.0) else "recomputed per chunk in the backward",
                       "l2_policy": "inputs (128 MB of points + 512 MB of jets/adjoints per step) exceed the 126 MB L2",
                       "collectives_per_step": 0 if world == 1 else 2},
            "phase_ms": {"forward_and_losses": fwd_ms, "backward": bwd_ms},
            "step_tflops_algorithmic": f_step * n / (ms_per_step * 1e-3) / 1e12,
            # the whole step against the FP32 FFMA roofline the reference's arithmetic (fp32) would be bound by: > 1 because
            # the contractions run as 3xTF32 on the tensor cores
            "step_frac_of_fp32_ffma_peak": f_step * n / (ms_per_step * 1e-3) / 1e12 / fp32_peak,
            "roofline": None,
            "roofline_ffma": {"bound": "fp32", "kernel": "jet_bwd_kernel<2,LAPLACIAN,128>", "achieved": bwd_alg, "peak": fp32_peak,
                         "unit": "TFLOP/s", "frac": bwd_alg / fp32_peak,
                         "peak_source": "FFMA/FFMA2 probe kernel timed in this run (MEASURED_PEAKS.json has no fp32 figure); "
                                        f"ffma={peak['ffma']:.1f} ffma2={peak['ffma2']:.1f} TFLOP/s",
                         "algorithmic_flops_per_point": 2 * f_fwd, "executed_flops_per_point": 3 * f_fwd4,
                         "tensor_peak_tflops_bf16_measured": measured.get("bf16_tflops_sustained"),
                         "hbm_gbs_measured": measured.get("hbm_gbs"), "traffic": None,
                         "point_stream_gbs": (8 + 16 + 16 + 4) * n / (bwd_ms * 1e-3) / 1e9},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": X_host.numel() * 4 * world, "d2h_bytes_per_step": 4 * world,
                    "includes": "pinned-host -> device copy of every step's collocation block (double-buffered on a copy stream, overlapping the previous step), loss, backward, all-reduce, Adam step, loss read-back",
                    "last_loss": last},
            "gpu_launches": kernels_per_step(n, tensor, bool(tc_ms) and tc_ms.get("stash_forward", 1.0) == 0.0) * args.steps,
            "clocks": clocks, "loss": float(loss.detach()),
        }
        if tensor and tc_ms and max(tc_ms.values()) > 0:
            d, W, L = LAYERS[0], LAYERS[1], len(LAYERS) - 2
            f_contr = 2 * W * W * (1 + 2 * d) * (L - 1)          # one hidden-layer contraction sweep per point (algorithmic)
            kname = max(tc_ms, key=tc_ms.get)
            kms = tc_ms[kname]
            alg = {"stash_forward": f_contr, "reverse_sweep": f_contr, "dw_contraction": f_contr, "fused_backward": 2 * f_contr}[kname]
            sym = {"stash_forward": "tcp::jet_fwd2_tc_kernel<2,true,ACT_TANH>", "reverse_sweep": "tcp::jet_rev_tc_kernel<2,ACT_TANH>",
                   "dw_contraction": "tcp::dw_tc_kernel<2,ACT_TANH>", "fused_backward": "tcp::bwd_fused_tc_kernel<2,ACT_TANH>"}[kname]
            bf16 = measured.get("bf16_tflops_sustained") or 1400.0
            achieved = alg * n / (kms * 1e-3) / 1e12
            # image blocks of 2 KB per point and hidden matrix l = 1..L-1: st_l written by the stash forward, read by the
            # reverse sweep (all) and the dW contraction (l <= L-2); zbar_l written by the reverse sweep, read by dW
            hbm_bytes = {"stash_forward": 8 + 2048 * (L - 1), "reverse_sweep": 8 + 16 + 2 * 2048 * (L - 1),
                         "dw_contraction": 8 + 2048 * (L - 1) + 2048 * (L - 2),
                         "fused_backward": 8 + 16 + 2048 * (L - 1) + 2048 * (L - 2)}[kname]
            # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch (= one 262 144-point chunk) from the committed
            # `ncu --set full` captures (profiles/r1b_tc_*_ncu_full_raw.csv); algorithmic bytes for the same launch beside it
            ncu_traffic = {"stash_forward": 1.554e9, "reverse_sweep": 3.194e9, "dw_contraction": 2.690e9, "fused_backward": None}[kname]
            line["roofline"] = {
                "bound": "tensor", "kernel": sym, "achieved": achieved, "peak": bf16, "unit": "TFLOP/s", "frac": achieved / bf16,
                "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained" if measured.get("bf16_tflops_sustained") else "fallback 1.4 PFLOP/s",
                "tf32_peak": bf16 / 2, "effective_3xtf32_peak": bf16 / 6, "frac_of_3xtf32_peak": achieved / (bf16 / 6),
                "algorithmic_flops_per_point": alg, "executed_tensor_flops_per_point": 3 * 2 * W * W * 4 * (L - 1),
                "kernel_ms_per_step": kms, "kernel_launches_per_step": (n + TC_CHUNK - 1) // TC_CHUNK,
                "tensor_backward_ms": tc_ms, "traffic": ncu_traffic, "traffic_unit": "bytes per launch (262144 points), ncu",
                "algorithmic_bytes_per_launch": hbm_bytes * min(n, TC_CHUNK),
                "algorithmic_hbm_bytes_per_point": hbm_bytes, "hbm_gbs_achieved": hbm_bytes * n / (kms * 1e-3) / 1e9,
                "hbm_gbs_measured": measured.get("hbm_gbs"), "fp32_ffma_peak_tflops": fp32_peak}
            del line["roofline_ffma"]
        else:
            line["roofline"] = line.pop("roofline_ffma")
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            times = cpu_reference_steps(args.cpu_points, 3, 1, threads)
            v = args.cpu_points / (sum(times) / len(times))
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": f"{args.cpu_points} of the 16M points per step x 3 steps, oracle/pinn_oracle.GPE2D "
                                              "(restated reference class) on the host cores"}
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
    ap.add_argument("--points", type=int, default=16 * 1024 * 1024, help="collocation points per GPU")
    ap.add_argument("--cpu-points", type=int, default=50000, help="bounded sample for the CPU port")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
