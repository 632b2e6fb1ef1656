#!/usr/bin/env python
"""SASS size of every kernel in the built objects (instruction-cache footprint; no GPU needed)."""
import re
import subprocess
import sys
from pathlib import Path

build = Path(__file__).resolve().parent.parent / "vibevoice_b200" / "csrc" / "build"
for obj in sorted(build.glob("vv_*.o")):
    out = subprocess.run(["cuobjdump", "-sass", str(obj)], capture_output=True, text=True).stdout
    name, last, sizes = None, 0, {}
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", line)
        if m and name:
            sizes[name] = int(m.group(1), 16) + 16
    for n, s in sorted(sizes.items(), key=lambda kv: -kv[1]):
        d = subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
        d = re.sub(r"\(anonymous namespace\)::", "", d)
        print(f"{s / 1024:8.1f} KB  {obj.name:14s} {d[:100]}")
