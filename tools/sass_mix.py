#!/usr/bin/env python
"""Per-kernel SASS summary of libb200fhe.so (cuobjdump -sass): instruction count and mnemonic histogram, the evidence behind
the instruction-level statements in DESIGN.md (IMAD.WIDE per butterfly / per MAC, UBLKCP = bulk TMA, no tensor-core ops).
usage: tools/sass_mix.py [substring ...] > profiles/rNN_sass_mix.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "phantom-fhe-boot_b200", "libb200fhe.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
filt = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", out)), capture_output=True, text=True).stdout.splitlines()
names = dict(zip(re.findall(r"Function : (\S+)", out), filt))
want = sys.argv[1:]
cur, hist = None, {}
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = names.get(m.group(1), m.group(1))
        hist[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1).split(".")
        hist[cur][".".join(op[:2]) if op[0] in ("IMAD", "UBLKCP", "LDG", "STG", "LDS", "STS", "SYNCS", "UTMALDG", "MEMBAR") else op[0]] += 1
for k in sorted(hist):
    short = re.sub(r"\(.*", "", k.replace("(anonymous namespace)::", "")).replace("b200fhe::", "")
    if want and not any(w in short for w in want):
        continue
    h = hist[k]
    n = sum(h.values())
    print("%-60s %5d instr | %s" % (short[:60], n, ", ".join("%s %d" % kv for kv in h.most_common(12))))
