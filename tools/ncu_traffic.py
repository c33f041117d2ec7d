#!/usr/bin/env python3
"""Summarise an `ncu --set full` capture (raw CSV page) of the dominant kernel into profiles/traffic.json.

usage: ncu -i capture.ncu-rep --page raw --csv > raw.csv
       tools/ncu_traffic.py raw.csv wf_extend "<command that was profiled>" [bench line .json of the same command, run without ncu]
bench.py reads the result for roofline.traffic / roofline.limiter; `kernel_source_hash` lets it flag a capture of older
kernels, `algorithmic_bytes_per_launch` (from the plain run's JSON line) lets it scale the DRAM bytes to other slices."""
import csv, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
raw, kernel, cmd = sys.argv[1], sys.argv[2], (sys.argv[3] if len(sys.argv) > 3 else "")
rows = list(csv.reader(open(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}


def val(r, name, scale_bytes=False):
    v = float(r[col[name]].replace(",", ""))
    return v * UNIT.get(units[col[name]], 1) if scale_bytes else v


sel = [r for r in data if kernel in r[col["Kernel Name"]]]
if not sel:
    raise SystemExit("no launch of " + kernel)
per = []
for r in sel:
    per.append({"ms": val(r, "gpu__time_duration.sum") * {"ms": 1, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(units[col["gpu__time_duration.sum"]], 1),
                "dram_read": val(r, "dram__bytes_read.sum", True), "dram_write": val(r, "dram__bytes_write.sum", True)})
keep = ["launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct"]
n = len(per)
out = {
    "kernel": sel[0][col["Kernel Name"]],
    "command": cmd,
    "launches_captured": n,
    "kernel_source_hash": bench.kernel_source_hash(),
    "algorithmic_bytes_per_launch": (json.load(open(sys.argv[4]))["roofline"]["algorithmic"]["bytes_per_launch"] if len(sys.argv) > 4 else None),
    "dram_bytes_per_launch": sum(p["dram_read"] + p["dram_write"] for p in per) / n,
    "dram_read_bytes_per_launch": sum(p["dram_read"] for p in per) / n,
    "dram_write_bytes_per_launch": sum(p["dram_write"] for p in per) / n,
    "ms_per_launch_under_ncu": sum(p["ms"] for p in per) / n,
    "per_launch": per if n <= 12 else per[:12],
    "metrics_mean": {k: sum(val(r, k) for r in sel) / n for k in keep if k in col},
}
dst = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "traffic.json")
json.dump(out, open(dst, "w"), indent=1)
print(json.dumps({k: v for k, v in out.items() if k != "per_launch"}, indent=1))
