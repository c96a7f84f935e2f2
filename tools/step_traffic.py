#!/usr/bin/env python
"""Sums an ncu CSV (metrics dram__bytes_read.sum, dram__bytes_write.sum, gpu__time_duration.sum) over the kernels of ONE training
step: the launches between the (skip+1)-th and (skip+2)-th occurrence of the marker kernel.  Usage:
    python tools/step_traffic.py raw.csv [marker=sample_depths_kernel] [skip=3]  >  summary.json"""
import csv
import json
import sys


def main(path, marker="sample_depths_kernel", skip=3):
    lines = [l for l in open(path, newline="") if not l.startswith("==")]
    per = {}
    order = []
    for r in csv.DictReader(lines):
        kid = r["ID"]
        if kid not in per:
            per[kid] = {"name": r["Kernel Name"].split("(")[0]}
            order.append(kid)
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "")
        if r["Metric Name"].startswith("dram__bytes"):
            v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
        elif r["Metric Name"].startswith("gpu__time"):
            v *= {"ns": 1e-3, "nsecond": 1e-3, "us": 1, "usecond": 1, "ms": 1e3, "msecond": 1e3}.get(unit, 1e-3)
        per[kid][r["Metric Name"]] = v
    marks = [i for i, k in enumerate(order) if marker in per[k]["name"]]
    a, b = marks[skip], marks[skip + 1]
    ks = [per[k] for k in order[a:b]]
    rd = sum(k.get("dram__bytes_read.sum", 0) for k in ks)
    wr = sum(k.get("dram__bytes_write.sum", 0) for k in ks)
    us = sum(k.get("gpu__time_duration.sum", 0) for k in ks)
    by = {}
    for k in ks:
        e = by.setdefault(k["name"], [0, 0.0, 0.0])
        e[0] += 1
        e[1] += k.get("dram__bytes_read.sum", 0) + k.get("dram__bytes_write.sum", 0)
        e[2] += k.get("gpu__time_duration.sum", 0)
    top = sorted(by.items(), key=lambda kv: -kv[1][1])[:12]
    print(json.dumps({"kernels_in_step": len(ks), "dram_read_MB": round(rd / 1e6, 1), "dram_write_MB": round(wr / 1e6, 1),
                      "dram_total_MB": round((rd + wr) / 1e6, 1), "serialised_kernel_time_us": round(us, 1),
                      "top_by_traffic": [{"kernel": n, "launches": c, "MB": round(bts / 1e6, 1), "us": round(t, 1)} for n, (c, bts, t) in top]}, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], *(sys.argv[2:3] or ["sample_depths_kernel"]), *(int(x) for x in sys.argv[3:4]))
