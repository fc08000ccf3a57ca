"""Per-launch DRAM traffic of the captured kernels from an `ncu --set full` raw-page CSV
(ncu -i X.ncu-rep --page raw --csv > X.csv):  python tools/ncu_traffic.py X.csv out.json [tag]
out.json is what bench.py reads to fill roofline.traffic (bytes per launch, averaged over the
captured launches of the dominant kernel)."""
import csv
import json
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    rd, wr, nm, gr, du = (col[k] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum",
                                            "Kernel Name", "Grid Size", "gpu__time_duration.sum"))
    per = []
    for r in data:
        per.append({"kernel": r[nm].split("(")[0].replace("void ", ""), "grid": r[gr],
                    "dram_read_bytes": float(r[rd]) * UNIT[units[rd]],
                    "dram_write_bytes": float(r[wr]) * UNIT[units[wr]],
                    "time_us_under_ncu": float(r[du])})
    name = max(set(p["kernel"] for p in per), key=lambda k: sum(p["kernel"] == k for p in per))
    sel = [p for p in per if p["kernel"] == name]
    avg = sum(p["dram_read_bytes"] + p["dram_write_bytes"] for p in sel) / len(sel)
    out = {"tag": sys.argv[3] if len(sys.argv) > 3 else "", "kernel": name, "launches_captured": len(sel),
           "traffic_bytes_per_launch": avg,
           "note": "dram__bytes_read.sum + dram__bytes_write.sum, mean over the captured launches "
                   "(ncu --set full --clock-control none); outputs that stay in the 126 MB L2 are not "
                   "written back inside the kernel's window, so this can be below the algorithmic bytes",
           "per_launch": per}
    json.dump(out, open(sys.argv[2], "w"), indent=1)
    print(name, len(sel), "launches, mean DRAM traffic", avg / 1e6, "MB")


if __name__ == "__main__":
    main()
