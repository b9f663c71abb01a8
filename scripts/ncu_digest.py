"""Digest of one `ncu --set full` report: headline metrics + SASS opcode / stall histogram + hottest lines.
usage: ncu_digest.py report.ncu-rep [kernel-name-regex]   (a report that holds several kernels: one digest per regex)"""
import csv, collections, subprocess, sys, io
rep = sys.argv[1]
flt = ["--kernel-name", "regex:" + sys.argv[2]] if len(sys.argv) > 2 else []
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"] + flt, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_pipe_tc_wavefronts_mem_shared.sum", "smsp__inst_executed.sum", "sm__cycles_elapsed.max", "smsp__cycles_active.avg",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "sm__inst_executed_pipe_tensor.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread"]
for h, u, v in zip(hdr, units, vals):
    if h in want: print(f"{h:70s} {v} {u}")
print("kernel:", vals[hdr.index("Kernel Name")][:100])
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + flt, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
data = []
for r in rows[2:]:          # ncu prints the kernel's listing twice: keep the first block
    if len(r) != len(hdr): break
    data.append(r)
iS = hdr.index("# Samples"); iI = hdr.index("Instructions Executed"); iSrc = hdr.index("Source")
tot_s = sum(int(r[iS]) for r in data); tot_i = sum(int(r[iI]) for r in data)
print("total samples", tot_s, "warp instructions", tot_i)
op = collections.Counter(); ops = collections.Counter()
for r in data:
    o = [x for x in r[iSrc].split() if not x.startswith('@')][0].split('.')[0]
    op[o] += int(r[iI]); ops[o] += int(r[iS])
for k, v in op.most_common(16): print(f"  {k:10s} inst {100*v/tot_i:5.1f}%  samples {100*ops[k]/tot_s:5.1f}%")
stall = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {h: sum(int(r[hdr.index(h)] or 0) for r in data) for h in stall}
print("stalls:", ", ".join(f"{k[6:]} {100*v/tot_s:.0f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:7]))
print("hottest SASS lines (index, samples%, executed, text):")
for i, r in sorted(enumerate(data), key=lambda ir: -int(ir[1][iS]))[:18]: print(f"  {i:5d} {100*int(r[iS])/tot_s:5.1f}% {r[iI]:>9s} {r[iSrc][:80]}")
