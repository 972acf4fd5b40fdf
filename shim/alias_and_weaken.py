#!/usr/bin/env python
"""For every `mangled alias` line of renames.txt whose symbol is DEFINED in the object: add a global alias symbol at the
same address (the reference's compiled body stays reachable as `alias`) and make the original name weak, so that the shim's
strong definition wins for every reference to it — those inside the same reference object included (-fPIC code calls
interposable functions through their symbol).  The reference's sources are not touched; this edits a COPY of the object.
usage: alias_and_weaken.py renames.txt in.o out.o"""
import subprocess
import sys

renames, src, dst = sys.argv[1:4]
pairs = [l.split() for l in open(renames) if l.strip() and not l.startswith("#")]
table = subprocess.run(["objdump", "-t", src], capture_output=True, text=True, check=True).stdout.splitlines()
args = []
for mangled, alias in pairs:
    for line in table:
        f = line.split()
        # value flags... section size name ; global function definitions only
        if len(f) >= 6 and f[-1] == mangled and " g " in line and " F " in line and f[-3] != "*UND*":
            args += ["--add-symbol", "%s=%s:0x%s,global,function" % (alias, f[-3], f[0]), "--weaken-symbol", mangled]
            break
subprocess.run(["objcopy", *args, src, dst], check=True)
print("%s: %d symbol(s) aliased and weakened" % (dst, len(args) // 4))
