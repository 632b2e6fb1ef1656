#!/usr/bin/env python
"""Ceiling of a TMA weight stream: one (or two) CTAs per SM pull a buffer larger than L2 through a
TMA -> mbarrier ring and drop it (vv_debug_tma_stream).  Prints GB/s for box size x ring depth x
CTAs x copy flavour, plus the "every CTA reads the same slab" case (activation tiles)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from vibevoice_b200 import _lib  # noqa: E402
from vibevoice_b200._lib import ptr  # noqa: E402

lib = _lib.load()
nbytes = 1 << 30
buf = torch.empty(nbytes, dtype=torch.uint8, device="cuda").random_(0, 100)
st = torch.cuda.current_stream()
stp = C.c_void_p(st.cuda_stream)


def run(box_rows, stages, ctas, flavour, shared, use_bytes=nbytes):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for it in range(3):
        e0.record()
        rc = lib.vv_debug_tma_stream(ptr(buf), use_bytes, box_rows, stages, ctas, flavour, shared, stp)
        assert rc == 0, lib.vv_last_error()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    moved = use_bytes * (ctas if shared else 1)
    return moved / best / 1e6, best


print("private slabs (HBM stream, 1 GiB):  box KB, stages, CTAs, flavour -> GB/s (per-SM B/clk at 1.9 GHz)")
for flavour in (0, 1):
    for ctas in (148, 296):
        for box_rows in (64, 128, 256):
            for stages in (2, 4, 8, 16):
                if stages * box_rows * 128 * (2 if ctas > 148 else 1) > 200 * 1024:
                    continue
                gbs, ms = run(box_rows, stages, ctas, flavour, 0)
                print(f"  {'tensor' if flavour == 0 else 'bulk  '} ctas {ctas:3d} box {box_rows * 128 // 1024:2d} KB stages {stages:2d}: "
                      f"{gbs:7.0f} GB/s  ({gbs / 148 / 1.9:5.1f} B/clk/SM)  in flight/SM {stages * box_rows * 128 * (ctas // 148) // 1024} KB")
print("shared slab (every CTA reads the same 8 MiB, L2-resident): GB/s delivered to the SMs in total")
for ctas in (148, 64):
    for box_rows in (64, 128):
        for stages in (4, 8):
            gbs, ms = run(box_rows, stages, ctas, 0, 1, use_bytes=8 << 20)
            print(f"  tensor ctas {ctas:3d} box {box_rows * 128 // 1024:2d} KB stages {stages:2d}: {gbs:7.0f} GB/s "
                  f"({gbs / ctas / 1.9:5.1f} B/clk/SM)")
print("shared slab, box size sweep (stages 8 unless the ring would not fit): clk per box per SM")
for flavour in (0, 1):
    for ctas in (148, 296):
        for box_rows in (8, 16, 32, 64, 128, 256):
            stages = 8 if box_rows <= 64 or ctas == 148 and box_rows <= 128 else 4
            if ctas == 296 and box_rows == 256:
                stages = 2
            gbs, ms = run(box_rows, stages, ctas, flavour, 1, use_bytes=8 << 20)
            per_sm = gbs / 148 / 1.9
            print(f"  {'tensor' if flavour == 0 else 'bulk  '} ctas {ctas:3d} box {box_rows * 128 / 1024:4.0f} KB stages {stages}: "
                  f"{gbs:7.0f} GB/s ({per_sm:5.1f} B/clk/SM, {box_rows * 128 / per_sm * (ctas / 148):5.0f} clk between boxes of one CTA)")
print("GEMM-like traffic per k-block (16 KB shared activation box + 8 KB private weight box, one barrier): clk per k-block")
for ctas in (148, 128):
    for stages in (2, 4, 8):
        gbs, ms = run(192, stages, ctas, 2, 0)
        per_cta = (nbytes // 4) // (24 * 1024) // ctas
        print(f"  ctas {ctas} stages {stages}: {ms * 1.9e6 / per_cta:6.0f} clk per k-block, weights {per_cta * ctas * 8192 / ms / 1e6:6.0f} GB/s")
