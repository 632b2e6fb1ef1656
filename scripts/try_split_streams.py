#!/usr/bin/env python
"""Experiment: does splitting the batch over concurrent streams (one decoder handle per stream) raise the
throughput of the latency-bound streaming decode step?"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from vibevoice_b200 import _lib, synth  # noqa: E402
from vibevoice_b200._lib import check, ptr  # noqa: E402
from vibevoice_b200.decoder import AcousticDecoderB200  # noqa: E402

dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
lib = _lib.load()
dcfg = synth.DEC_FULL
W = synth.init_decoder(dcfg)
N = 64
for parts in (1, 2, 4):
    n = N // parts
    decs = [AcousticDecoderB200(W, dcfg, device=dev, max_rows=n) for _ in range(parts)]
    caches = [d.new_cache(n) for d in decs]
    streams = [torch.cuda.Stream() for _ in range(parts)]
    idx = torch.arange(n, device=dev)
    lat = [torch.randn(n, 64, device=dev).bfloat16() for _ in range(parts)]
    wave = [torch.empty(n, 1, dcfg.hop, device=dev, dtype=torch.bfloat16) for _ in range(parts)]

    def step():
        ev = torch.cuda.Event()
        ev.record()
        for i in range(parts):
            streams[i].wait_event(ev)
            check(lib.vv_dec_step(decs[i]._h, caches[i]._c, ptr(lat[i]), ptr(idx), n, 1, 1, 1.0, 0.0, ptr(wave[i]),
                                  _lib.VV_DTYPE_BF16, C.c_void_p(streams[i].cuda_stream)), "vv_dec_step")
        for i in range(parts):
            torch.cuda.current_stream().wait_stream(streams[i])

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        step()
    e1.record()
    torch.cuda.synchronize()
    print(f"decode step, {N} utterances as {parts} x {n} on {parts} stream(s): {e0.elapsed_time(e1) / 20:.4f} ms")
    del decs, caches
