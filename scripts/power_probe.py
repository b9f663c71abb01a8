"""Is a kernel clock-bound or power-bound?  Runs one tensor-path call in a loop for a couple of seconds while nvidia-smi
samples SM clock and board power every 100 ms, and reports ms per call, median MHz / W and the product (cycles per call).
Two schedules with the same ms per call but different cycles per call are limited by the power cap, not by the schedule.
usage: power_probe.py <fwd|stash|bwd> [n_points] [seconds]    (kernel variants are chosen through the GPE_* environment)"""
import os, sys, json, subprocess, threading, time, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import torch
from gpe_b200 import ops
from test_jet_kernels import make_net

what = sys.argv[1] if len(sys.argv) > 1 else "bwd"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 22
secs = float(sys.argv[3]) if len(sys.argv) > 3 else 3.0
d, W, L, jet = 2, 128, 4, 3
Ws, bs = make_net(d, W, L, seed=1, act=0)
Ws = [w.cuda() for w in Ws]; bs = [b.cuda() for b in bs]
x = torch.rand(n, d, device="cuda") * 4 - 2
ub = torch.randn(n, device="cuda"); gb = torch.randn(n, d, device="cuda"); hb = torch.randn(n, 1, device="cuda")
stash = ops.mlp_jet_fwd_stash(x, Ws, bs, 0, jet)[3] if what == "bwd" else None


def call():
    if what == "fwd":
        ops.mlp_jet_fwd(x, Ws, bs, 0, jet)
    elif what == "stash":
        ops.mlp_jet_fwd_stash(x, Ws, bs, 0, jet)
    else:
        ops.mlp_jet_bwd_stashed(x, Ws, bs, 0, jet, ub, gb, hb, stash)


for _ in range(3): call()
torch.cuda.synchronize()
samples = []
proc = subprocess.Popen(["nvidia-smi", "-i", "0", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.active", "--format=csv,noheader,nounits", "-lms", "100"],
                        stdout=subprocess.PIPE, text=True)


def reader():
    for line in proc.stdout:
        p = [s.strip() for s in line.split(",")]
        try: samples.append((time.time(), float(p[0]), float(p[1]), p[2]))
        except Exception: pass


threading.Thread(target=reader, daemon=True).start()
time.sleep(0.3)
t0 = time.time()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); calls = 0
while time.time() - t0 < secs:
    for _ in range(5): call()
    calls += 5
    torch.cuda.synchronize()
b.record(); torch.cuda.synchronize()
t1 = time.time()
proc.terminate()
busy = [s for s in samples if t0 + 0.5 <= s[0] <= t1]
ms = a.elapsed_time(b) / calls
mhz = statistics.median(s[1] for s in busy) if busy else float("nan")
watt = statistics.median(s[2] for s in busy) if busy else float("nan")
timings = None
if what == "bwd":
    ops.tc_profile(True); call(); timings = ops.tc_timings(); ops.tc_profile(False)
print(json.dumps({"what": what, "n": n, "env": {k: v for k, v in os.environ.items() if k.startswith("GPE_")}, "ms_per_call": round(ms, 3),
                  "sm_mhz": mhz, "watt": watt, "mcycles_per_call": round(ms * mhz / 1e3, 2), "samples": len(busy),
                  "reasons": sorted({s[3] for s in busy}), "kernels_ms": timings}), flush=True)
