#!/usr/bin/env python
"""Measurement aid: cost of tcgen05.commit for the issuing thread (vv_debug_mma_commit)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from vibevoice_b200 import _lib  # noqa: E402

torch.zeros(1, device="cuda")
lib = _lib.load()
out = (C.c_longlong * 148)()
print("bn mmas nacc -> clk per iteration for 0/1/2 commits (median over CTAs); nacc=0: 4 unrolled MMAs issued by one thread in divergent code; nacc=-1: converged warp + elect.sync")
for ctas in (148,):
    for bn in (32, 64, 128, 256):
        for mmas, nacc in ((4, 0), (4, -1), (8, -1), (16, -1)):
            if nacc * bn > 256:
                continue
            row = []
            for commits in (0, 1, 2):
                rc = lib.vv_debug_mma_commit(bn, mmas, commits, 2000, out, ctas, nacc)
                assert rc == 0, lib.vv_last_error()
                row.append(int(np.median(np.array(out[:ctas]))))
            print(f"ctas={ctas:3d} bn={bn:3d} mmas={mmas:2d} nacc={nacc}: {row}")

print("two issuing warps (nacc=-2), two warps + 2 complete waits per iteration (-3), one warp + 2 waits (-4): clk per iteration of EACH warp")
for bn in (64, 128):
    for mmas in (4, 8, 16):
        for nacc in (-4, -3, -2):
            row = []
            for commits in (0, 1, 2):
                rc = lib.vv_debug_mma_commit(bn, mmas, commits, 2000, out, 148, nacc)
                assert rc == 0, lib.vv_last_error()
                row.append(int(np.median(np.array(out[:148]))))
            print(f"bn={bn:3d} mmas={mmas:2d} nacc={nacc}: {row}")
