#!/usr/bin/env python
"""SASS size of every kernel in the built objects (instruction-cache footprint) and, with --mnemonics, the per-kernel
counts of the instructions that prove tcgen05 / TMEM / TMA use (no GPU needed):

    python scripts/sass_sizes.py                 # sizes
    python scripts/sass_sizes.py --mnemonics     # markdown table -> profiles/r02_sass_summary.md

UTCHMMA = tcgen05.mma, UTCBAR = tcgen05.commit, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG = TMA tensor load/store,
UTMAPF = TMA L2 prefetch, UBLKCP = cp.async.bulk (1-D), SYNCS = mbarrier, HMMA = mma.sync."""
import re
import subprocess
import sys
from pathlib import Path

MNEMONICS = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UBLKCP", "SYNCS", "HMMA"]
build = Path(__file__).resolve().parent.parent / "vibevoice_b200" / "csrc" / "build"
want_mn = "--mnemonics" in sys.argv
rows = []
for obj in sorted(build.glob("vv_*.o")):
    out = subprocess.run(["cuobjdump", "-sass", str(obj)], capture_output=True, text=True).stdout
    name, sizes, counts = None, {}, {}
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            counts[name] = dict.fromkeys(MNEMONICS, 0)
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*)", line)
        if m and name:
            sizes[name] = int(m.group(1), 16) + 16
            op = re.match(r"(@!?U?P\w+\s+)?([A-Z0-9_]+)", m.group(2))
            if op and op.group(2) in counts[name]:
                counts[name][op.group(2)] += 1
    for n, sz in sorted(sizes.items(), key=lambda kv: -kv[1]):
        d = subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
        d = re.sub(r"\(anonymous namespace\)::", "", d)
        d = re.sub(r"\(.*", "", d)
        rows.append((obj.name, d, sz, counts[n]))
if want_mn:
    print("| object | kernel | SASS KB | " + " | ".join(MNEMONICS) + " |")
    print("|---|---|---|" + "---|" * len(MNEMONICS))
    for o, d, sz, c in rows:
        if any(c[m] for m in MNEMONICS if m != "SYNCS") or sz > 8192:
            print(f"| {o} | `{d[:70]}` | {sz / 1024:.1f} | " + " | ".join(str(c[m]) if c[m] else "" for m in MNEMONICS) + " |")
else:
    for o, d, sz, c in rows:
        print(f"{sz / 1024:8.1f} KB  {o:14s} {d[:100]}")
