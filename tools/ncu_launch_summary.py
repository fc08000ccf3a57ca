"""Per-kernel totals / shares from an `ncu --metrics gpu__time_duration.sum --csv` launch list.
usage: python tools/ncu_launch_summary.py gpurun_out/launches_X.csv"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
hdr, data = rows[hi], rows[hi + 1:]
ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in data:
    if len(r) <= iv:
        continue
    scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[iu], 1e-3)
    name = r[ik].split("(")[0][:64]
    agg[name][0] += 1
    agg[name][1] += float(r[iv].replace(",", "")) * scale
tot = sum(v[1] for v in agg.values())
print(f"{len(data)} launches, {tot:.1f} us total (cold-cache, serialised under ncu: compare shares)")
for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{k:66s} n={v[0]:4d} total={v[1]:9.1f} us avg={v[1] / v[0]:7.2f} us share={v[1] / tot:.3f}")
