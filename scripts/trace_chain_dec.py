#!/usr/bin/env python
"""Trace build only: per-phase clock stamps of the decoder's stage-0 program (k_chain, 24 phases:
8 x [mixer row phase | FFN1 + GELU | FFN2 split-K]) during one streaming frame-step at n utterances."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from vibevoice_b200 import _lib, synth  # noqa: E402
from vibevoice_b200._lib import check, ptr  # noqa: E402
from vibevoice_b200.decoder import AcousticDecoderB200  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
lib = _lib.load()
dev = torch.device("cuda:0")
dcfg = synth.DEC_FULL
dec = AcousticDecoderB200(synth.init_decoder(dcfg), dcfg, device=dev, max_rows=n)
cache = dec.new_cache(n)
idx = torch.arange(n, device=dev)
stp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
lat = torch.randn(n, dcfg.vae_dim, device=dev).bfloat16()
wave = torch.empty(n, 1, dcfg.hop, device=dev, dtype=torch.bfloat16)
for _ in range(4):
    check(lib.vv_dec_step(dec._h, cache._c, ptr(lat), ptr(idx), n, 1, 1, 1.0, 0.0, ptr(wave), _lib.VV_DTYPE_BF16, stp), "vv_dec_step")
torch.cuda.synchronize()
lib.vv_set_option(b"trace_select", 2)
NS = 384
N2 = 160 * 24 * 3 * 16
buf = (C.c_longlong * (160 * NS + N2))()
k = lib.vv_debug_trace(buf, 160 * NS + N2)
assert k > 0, "not a trace build"
names = sum([["mix", "ffn1", "ffn2"] for _ in range(8)], [])
t = np.array(buf[:160 * NS]).reshape(160, NS)[:148].reshape(148, 24, 16)
t2 = np.array(buf[160 * NS:160 * NS + N2]).reshape(160, 24, 3, 16)[:148]


def st(v):
    return f"{int(v.min()):6d} {int(np.median(v)):6d} {int(v.max()):6d}"


print("per phase over CTAs, min/med/max clk: work = done - prev exit | poll = arrival -> barrier seen complete")
for ph in range(1, 24):
    r = t[:, ph]
    work = r[:, 0] - t[:, ph - 1, 4]
    print(f" {ph:2d} {names[ph]:5s} work {st(work)} | poll {st(r[:, 3] - r[:, 2])} | exit {st(r[:, 4] - r[:, 3])}")
g = t[:, :, 7]
print("phase length (ns, median CTA):", [int(np.median(g[:, ph] - g[:, ph - 1])) for ph in range(1, 24)])
print("per ring group, clk since the phase started (median over CTAs that had an item)")
for ph in (4, 5, 7, 8):
    start = t[:, ph - 1, 4]
    sel = (t[:, ph, 0] - start) > 1500
    ng = 16
    for what, label in ((2, "W issued"), (0, "W landed"), (1, "A landed")):
        print(f" {ph:2d} {names[ph]:5s} {label}:", [int(np.median(t2[sel, ph, what, j] - start[sel])) for j in range(ng)])
    print(f" {ph:2d} {names[ph]:5s} epilogue:", [int(np.median(t2[sel, ph, 0, j] - start[sel])) for j in (12, 13, 14, 15)], "done",
          int(np.median(t[sel, ph, 0] - start[sel])), "active CTAs", int(sel.sum()))
