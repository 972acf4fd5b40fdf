#!/usr/bin/env python
"""Digest of one `ncu --set full` report: the handful of raw metrics the roofline discussion needs."""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "sm__inst_executed.sum", "sm__inst_issued.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warp_latency_issue_stalled_math_pipe_throttle_per_warp_active.pct" ]


def digest(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    u = dict(zip(hdr, units))
    all_ = []
    for vals in rows[2:]:   # one row per profiled launch
        if len(vals) != len(hdr):
            continue
        d = dict(zip(hdr, vals))
        res = {}
        for k in hdr:
            if k in KEYS or "issue_stalled" in k and k.endswith("_per_warp_active.pct") or k.startswith("smsp__pcsamp_warps_issue_stalled") and not k.endswith("not_issued"):
                res[k] = (d[k], u[k])
        all_.append((d, res))
    return all_


if __name__ == "__main__":
    for p in sys.argv[1:]:
      for d, res in digest(p):
        print("==", p, d.get("Kernel Name"), d.get("Grid Size"), d.get("Block Size"))
        stalls = []
        for k, (v, un) in res.items():
            if "pcsamp" in k:
                try:
                    stalls.append((float(v.replace(",", "")), k.replace("smsp__pcsamp_warps_issue_stalled_", "")))
                except ValueError:
                    pass
            elif "issue_stalled" in k:
                continue
            else:
                print("  %-78s %s %s" % (k, v, un))
        tot = sum(s for s, _ in stalls) or 1
        print("  stall samples:", ", ".join("%s %.0f%%" % (n, 100 * s / tot) for s, n in sorted(stalls, reverse=True)[:7]))
