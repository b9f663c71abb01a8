"""Summarise an `ncu --csv` launch list: per kernel, mean of each metric per launch."""
import csv, collections, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value"); mi = hdr.index("Metric Name")
agg = collections.defaultdict(lambda: collections.defaultdict(float)); cnt = collections.Counter()
for r in rows[1:]:
    try: v = float(r[vi].replace(",", ""))
    except ValueError: continue
    k = r[ki][:48]; agg[k][r[mi]] += v
    if r[mi] == "gpu__time_duration.sum": cnt[k] += 1
tot = sum(m["gpu__time_duration.sum"] for m in agg.values())
for k, m in sorted(agg.items(), key=lambda kv: -kv[1]["gpu__time_duration.sum"])[:10]:
    c = max(cnt[k], 1)
    extra = " ".join(f"{name.split('.')[0][-22:]}={val / c:.3g}" for name, val in m.items() if name != "gpu__time_duration.sum")
    print(f"{m['gpu__time_duration.sum'] / c / 1e6:8.3f} ms/launch x{c:3d} share {100 * m['gpu__time_duration.sum'] / tot:5.1f}%  {k}  {extra}")
