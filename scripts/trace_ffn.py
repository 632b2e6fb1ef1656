#!/usr/bin/env python
"""Trace build only (make -C vibevoice_b200/csrc TRACE=1): clock stamps of the fused FFN kernel (C = 256)
for the first tile of every CTA, last stage-3 block of one streaming frame-step."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from vibevoice_b200 import _lib, synth  # noqa: E402
from vibevoice_b200.decoder import AcousticDecoderB200  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
lib = _lib.load()
cfg = synth.DEC_FULL
dec = AcousticDecoderB200(synth.init_decoder(cfg), cfg, max_rows=n)
cache = dec.new_cache(n)
idx = torch.arange(n, device="cuda")
lat = torch.randn(n, 1, 64, device="cuda").bfloat16()
for _ in range(3):
    dec.decode(lat, cache=cache, sample_indices=idx, use_cache=True)
torch.cuda.synchronize()
lib.vv_set_option(b"trace_select", 3)
buf = (C.c_longlong * (160 * 48))()
k = lib.vv_debug_trace(buf, 160 * 48)
if k > 0:
    t = np.array(buf[:k]).reshape(160, 48)
    for cta in (0, 50, 99):
        r = t[cta]
        base = r[46]
        print(f"cta {cta}: per chunk (clk since the compute warps were ready): d1_full, h_free, epilogue-1 done")
        for c in range(8):
            print(f"   chunk {c}: {r[4*c]-base:7d} {r[4*c+1]-base:7d} {r[4*c+2]-base:7d}")
        for c in range(4, 8):
            print(f"   chunk {c} inside epilogue 1: h_free {r[4*c+1]-base}  tmem loaded {r[32+(c-4)*2]-base}  gelu+stores done {r[33+(c-4)*2]-base}  fenced+arrived {r[4*c+2]-base}")
        print(f"   d2_full {r[40]-base}  epilogue-2 done {r[41]-base}")
