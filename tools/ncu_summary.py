#!/usr/bin/env python
"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list (the recipe of B200_PROFILING.md) per kernel:
launches, total and mean microseconds, share of the listed time.  Usage: python tools/ncu_summary.py launches.csv [> summary.csv]
(`profiles/*_summary.csv` are made with this script from the raw lists beside them.)"""
import csv
import sys
from collections import OrderedDict


def main(path):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    rd = csv.DictReader(lines)
    for r in rd:
        if r.get("Metric Name", "") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        us = v / 1e3 if unit in ("ns", "nsecond") else v * 1e3 if unit in ("ms", "msecond") else v
        rows.append((r["Kernel Name"], us, r.get("Stream", "")))
    tot = sum(u for _, u, _ in rows) or 1.0
    agg = OrderedDict()
    for name, us, st in rows:
        short = name.split("(")[0]
        a = agg.setdefault(short, [0, 0.0, set()])
        a[0] += 1
        a[1] += us
        a[2].add(st)
    print("kernel,launches,total_us,mean_us,share,cuda_streams")
    for name, (n, us, sts) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%s,%d,%.1f,%.2f,%.4f,%d" % (name.replace(",", ";"), n, us, us / n, us / tot, len(sts)))


if __name__ == "__main__":
    main(sys.argv[1])
