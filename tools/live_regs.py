#!/usr/bin/env python
"""Per kernel: live GPR count at the first instruction and the maximum, from `nvdisasm -plr`.
A large count at entry means some variable is read on a path where it was never written (it then stays live
across the whole kernel and inflates the register allocation). usage: tools/live_regs.py file.cubin"""
import re
import subprocess
import sys

out = subprocess.run(["nvdisasm", "-plr", sys.argv[1]], stdout=subprocess.PIPE, text=True).stdout
name, first, mx = None, None, 0
for line in out.splitlines():
    m = re.match(r"\.text\.(\S+):", line)
    if m:
        if name:
            print(f"{name[:90]:90s} entry {first} max {mx}")
        name, first, mx = m.group(1), None, 0
        continue
    m = re.search(r"/\*[0-9a-f]{4,}\*/.*// \|\s*(\d+)", line)
    if m and name:
        n = int(m.group(1))
        if first is None:
            first = n
        mx = max(mx, n)
if name:
    print(f"{name[:90]:90s} entry {first} max {mx}")
