"""Write profiles/r2_ncu_traffic.json -- what bench.py reports as `roofline.traffic`: dram__bytes_read.sum +
dram__bytes_write.sum per launch of the tensor-path kernels, read from `ncu --set full` reports, with the git sha of the
tree that was profiled.  usage: ncu_traffic.py <points_per_launch>:report.ncu-rep [<points>:report2.ncu-rep ...]"""
import csv, io, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = {"jet_fwd2_tc_kernel": "stash_forward", "jet_fwd_tc_kernel": "stash_forward", "jet_rev_tc_kernel": "reverse_sweep",
        "jet_rev2_tc_kernel": "reverse_sweep", "dw_tc_kernel": "dw_contraction", "bwd_fused_tc_kernel": "fused_backward"}
out = {"git_sha": subprocess.run(["git", "-C", ROOT, "rev-parse", "HEAD"], capture_output=True, text=True).stdout.strip(),
       "how": "ncu --set full --clock-control none, one launch per kernel, scripts/tc_kernel_times.py", "kernels": {}}
def num(v, unit):
    return float(v) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}.get(unit, 1.0)
for arg in sys.argv[1:]:
    points, rep = arg.split(":", 1)
    points = int(points)
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        key = next((v for k, v in KEYS.items() if k + "<" in name), None)
        if key is None:
            continue
        rd = num(r[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]])
        wr = num(r[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]])
        dur = float(r[col["gpu__time_duration.sum"]]) * {"us": 1.0, "ms": 1e3, "ns": 1e-3}.get(units[col["gpu__time_duration.sum"]], 1.0)
        out["kernels"][key] = {"kernel": name.split("(")[0][:80], "points_per_launch": points, "dram_bytes_read": rd, "dram_bytes_write": wr,
                               "dram_bytes_per_launch": rd + wr, "dram_bytes_per_point": (rd + wr) / points, "duration_us_under_ncu": dur,
                               "tensor_pipe_active_pct": float(r[col["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"]]),
                               "report": os.path.basename(rep)}
json.dump(out, open(os.path.join(ROOT, "profiles", "r2_ncu_traffic.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
