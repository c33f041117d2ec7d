#!/usr/bin/env python3
"""Hot source lines of one kernel.
usage: ncu -i X.ncu-rep --page source --print-source cuda,sass --csv --kernel-name regex:K --launch-count 1 > k.csv
       tools/ncu_hot.py k.csv [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
data, fname = [], ""
hdr = None
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif len(r) > 8 and r[0] == "Line No":
        hdr = r
        ni, ii, ti = hdr.index('# Samples'), hdr.index('Instructions Executed'), hdr.index('Thread Instructions Executed')
    elif hdr and len(r) > 8 and r[0] != "":
        try:
            data.append((int(r[ni] or 0), int(r[ii] or 0), int(r[ti] or 0), fname, r[0], r[1]))
        except ValueError:
            pass
tot_s, tot_i = sum(d[0] for d in data), sum(d[1] for d in data)
print('total samples', tot_s, 'warp instructions', tot_i)
for d in sorted(data, key=lambda d: -d[0])[:top]:
    print(f"{d[0]:6d} {d[0] / max(tot_s, 1):5.1%} | inst {d[1]:10d} {d[1] / max(tot_i, 1):5.1%} | lanes {d[2] / max(d[1], 1):4.1f} | {d[3]}:{d[4]:4s} {d[5].strip()[:100]}")
