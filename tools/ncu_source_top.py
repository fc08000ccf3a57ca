"""Top stalled SASS instructions of one kernel from `ncu --page source --csv` output.
usage: ncu -i rep --page source --csv --kernel-name regex:X --launch-count 1 > src.csv; python tools/ncu_source_top.py src.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
data = []
for r in rows[hi + 1:]:
    if r and r[0] == "Kernel Name":
        break                      # first kernel block only
    if len(r) == len(hdr) and r[0].startswith("0x"):
        data.append(r)
isrc, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
tot = sum(int(r[isamp]) for r in data)
print("kernel:", rows[0][1][:100] if rows[0] else "?")
print("total samples", tot, "instructions", len(data))
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
agg = {}
for r in data:
    for i in stall_cols:
        agg[hdr[i][6:]] = agg.get(hdr[i][6:], 0) + int(r[i])
print("stall totals:", sorted(agg.items(), key=lambda x: -x[1])[:8])
top = sorted(enumerate(data), key=lambda x: -int(x[1][isamp]))[:n]
for idx, r in sorted(top):
    st = sorted(((hdr[i][6:], int(r[i])) for i in stall_cols if int(r[i]) > 0), key=lambda x: -x[1])[:3]
    print(f"{idx:5d} {int(r[isamp]):6d} {100*int(r[isamp])/max(tot,1):5.1f}% ex={r[iex]:>8s} {r[isrc].strip()[:64]:64s} {st}")
