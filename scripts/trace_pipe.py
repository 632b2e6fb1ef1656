#!/usr/bin/env python
"""Trace build only (make -C vibevoice_b200/csrc TRACE=1 BUILD=build_trace OUT=../libvvb200_trace.so, run with
VVB200_LIB=vibevoice_b200/libvvb200_trace.so): per-tile clock stamps of the pipelined Block1D kernel
(k_block_pipe) for the last decoder block of one streaming frame-step.
    python scripts/trace_pipe.py [n] [stages]     stages 7 -> C = 32, 6 -> C = 64, 5 -> C = 128
"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from vibevoice_b200 import _lib, synth  # noqa: E402
from vibevoice_b200.decoder import AcousticDecoderB200  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
stages = int(sys.argv[2]) if len(sys.argv) > 2 else 7
lib = _lib.load()
full = synth.DEC_FULL
cfg = synth.DecCfg(vae_dim=64, n_filters=32 * 2 ** (7 - stages), ratios=full.ratios[:stages - 1], depths=full.depths[:stages])
dec = AcousticDecoderB200(synth.init_decoder(cfg), cfg, max_rows=n)
cache = dec.new_cache(n)
idx = torch.arange(n, device="cuda")
lat = torch.randn(n, 1, 64, device="cuda").bfloat16()
for _ in range(3):
    dec.decode(lat, cache=cache, sample_indices=idx, use_cache=True)
torch.cuda.synchronize()
lib.vv_set_option(b"trace_select", 4)
buf = (C.c_longlong * (160 * 48))()
k = lib.vv_debug_trace(buf, 160 * 48)
print("last-stage channels", cfg.channels(stages - 1), "trace elems", k)
if k > 0:
    t = np.array(buf[:k]).reshape(160, 48)
    for cta in (0, 77, 147):
        r = t[cta]
        base = r[4]
        for it in range(5):
            e = r[it * 8:it * 8 + 4] - base      # epilogue warp: y_full seen, first D1 ready, gelu done, epilogue 2 of the PREVIOUS tile done
            m = r[it * 8 + 4:it * 8 + 8] - base  # mixer: tile start, xraw landed, y_free passed (norm done), conv done
            print(f"cta {cta} tile {it}: MIX start {m[0]} xraw +{m[1]-m[0]} norm(+y_free) +{m[2]-m[1]} conv +{m[3]-m[2]} (total {m[3]-m[0]})"
                  f" | EPI y_full@{e[0]} d1wait +{e[1]-e[0]} gelu +{e[2]-e[1]} epi2(prev) +{e[3]-e[2]} (done@{e[3]})")
        for j in range(2):
            m = r[40 + j * 4:44 + j * 4] - base
            print(f"cta {cta} tile {j+1}: MMA warp: arrives at y_full wait @{m[0]}, passes @{m[1]}, last h_ready passed @{m[2]}, last mma2 issued @{m[3]}")
