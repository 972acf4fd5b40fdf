#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total time, share."""
import collections
import csv
import json
import re
import sys


def summarise(path):
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    tot = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "")
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1.0)
        tot[name][0] += 1
        tot[name][1] += v
    total = sum(v[1] for v in tot.values())
    rows = [{"kernel": k, "launches": v[0], "total_us": round(v[1], 1), "share": round(v[1] / total, 4),
             "avg_us": round(v[1] / v[0], 2)} for k, v in sorted(tot.items(), key=lambda kv: -kv[1][1])]
    return {"total_kernel_ms": round(total / 1e3, 3), "launches": sum(v[0] for v in tot.values()), "kernels": rows}


if __name__ == "__main__":
    r = summarise(sys.argv[1])
    if len(sys.argv) > 2:
        json.dump(r, open(sys.argv[2], "w"), indent=1)
    print("total %.3f ms over %d launches" % (r["total_kernel_ms"], r["launches"]))
    for k in r["kernels"]:
        print("%-34s n=%5d %9.1f us %5.1f%% avg %8.2f us" % (k["kernel"][:34], k["launches"], k["total_us"], 100 * k["share"], k["avg_us"]))
