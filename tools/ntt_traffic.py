#!/usr/bin/env python
"""profiles/rNN_ntt_traffic.json from one `ncu --set full` report of `bench.py --no-extra`: per launch of the headline
transform's kernels (grid (32, 36, 16) = 576 limbs) duration, DRAM bytes and pipe utilisation; the forward pair's DRAM bytes
are what bench.py reports as roofline.traffic.  usage: ntt_traffic.py report.ncu-rep out.json"""
import csv
import io
import json
import subprocess
import sys


def main(rep, out):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr = rows[0]

    def num(r, k):
        try:
            return float(r[hdr.index(k)].replace(",", ""))
        except (ValueError, IndexError):
            return None
    units = dict(zip(hdr, rows[1]))

    def to_bytes(v, k):
        u = units.get(k, "")
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    seen = {}
    for r in rows[2:]:
        if len(r) != len(hdr):
            continue
        name, grid = r[hdr.index("Kernel Name")], r[hdr.index("Grid Size")]
        if "36, 16" not in grid:
            continue
        key = name.split("(")[0]
        if key in seen:
            continue
        seen[key] = {"kernel": key, "grid": grid, "duration_us": num(r, "gpu__time_duration.sum"),
                     "dram_read_bytes": to_bytes(num(r, "dram__bytes_read.sum"), "dram__bytes_read.sum"),
                     "dram_write_bytes": to_bytes(num(r, "dram__bytes_write.sum"), "dram__bytes_write.sum"),
                     "fmaheavy_pct": num(r, "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed"),
                     "alu_pct": num(r, "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
                     "issue_pct": num(r, "sm__inst_issued.avg.pct_of_peak_sustained_active")}
    fwd = [v for k, v in seen.items() if "ntt_fwd" in k]
    inv = [v for k, v in seen.items() if "ntt_inv" in k]
    res = {"command": "ncu --set full --clock-control none -k regex:ntt_ python bench.py --no-extra --steps 2 --warmup 3",
           "fwd": {"kernels": fwd, "pair_dram_bytes": sum(k["dram_read_bytes"] + k["dram_write_bytes"] for k in fwd)},
           "inv": {"kernels": inv, "pair_dram_bytes": sum(k["dram_read_bytes"] + k["dram_write_bytes"] for k in inv)}}
    json.dump(res, open(out, "w"), indent=1)
    print(json.dumps(res)[:600])


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
