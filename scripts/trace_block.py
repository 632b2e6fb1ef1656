#!/usr/bin/env python
"""Trace build only (make -C vibevoice_b200/csrc TRACE=1): per-tile clock stamps of the fused
Block1D kernel for the last decoder block (C = 32) of one streaming frame-step."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from vibevoice_b200 import _lib, synth  # noqa: E402
from vibevoice_b200.decoder import AcousticDecoderB200  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
stages = int(sys.argv[2]) if len(sys.argv) > 2 else 7     # decode with fewer stages to trace C=64/128
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
lib.vv_set_option(b"trace_select", 1)
buf = (C.c_longlong * (160 * 48))()
k = lib.vv_debug_trace(buf, 160 * 48)
print("last-stage channels", cfg.channels(stages - 1), "trace elems", k)
if k > 0:
    t = np.array(buf[:k]).reshape(160, 48)
    for cta in (0, 77, 147):
        r = t[cta]
        base = r[0]
        for it in range(5):
            s = r[it * 8:it * 8 + 6] - base
            print(f"cta {cta} tile {it}: start {s[0]}  mixer_done +{s[1]-s[0]}  d1_full +{s[2]-s[1]}  gelu_done +{s[3]-s[2]}  "
                  f"d2_full +{s[4]-s[3]}  epi2_done +{s[5]-s[4]}  | tile total {s[5]-s[0]}")
