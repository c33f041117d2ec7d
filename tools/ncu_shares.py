#!/usr/bin/env python3
"""Per-kernel shares of an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv --log-file X.csv ...`).

usage: tools/ncu_shares.py launches.csv "<command that was profiled>" > profiles/rNN_launch_shares.json
ncu serialises kernels, so the shares are of the serial sum (compare with roofline.serial_step of bench.py)."""
import csv, json, re, sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
k, m, v, u = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}
out = {}
for r in rows[1:]:
    if r[m] != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*", "", r[k]).strip()
    d = out.setdefault(name, {"launches": 0, "ms": 0.0})
    d["launches"] += 1
    d["ms"] += float(r[v].replace(",", "")) * scale[r[u]]
total = sum(d["ms"] for d in out.values())
for d in out.values():
    d["share"] = d["ms"] / total
print(json.dumps({"command": sys.argv[2] if len(sys.argv) > 2 else "", "note": "ncu serialises kernels: shares are of the serial sum",
                  "total_ms": total, "kernels": dict(sorted(out.items(), key=lambda kv: -kv[1]["ms"]))}, indent=1))
