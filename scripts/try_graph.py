#!/usr/bin/env python
"""Config 5 with the frame-step captured in a CUDA graph (torch.cuda.CUDAGraph around the two public calls)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from vibevoice_b200 import synth  # noqa: E402
from vibevoice_b200.decoder import AcousticDecoderB200  # noqa: E402
from vibevoice_b200.head import DiffusionHeadB200  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1
dev = torch.device("cuda:0")
hcfg, dcfg = synth.HEAD_1P5B, synth.DEC_FULL
head = DiffusionHeadB200(synth.init_head(hcfg, seed=1234), hcfg, device=dev, max_rows=2 * n)
dec = AcousticDecoderB200(synth.init_decoder(dcfg, seed=1234), dcfg, device=dev, max_rows=n)
head.set_schedule(10, "dpmsolver++")
cache = dec.new_cache(n)
idx = torch.arange(n, dtype=torch.int64, device=dev)
cond, neg, noise = [t.to(dev) for t in synth.synth_inputs(n, hcfg.hidden, hcfg.latent)]
cond, neg = cond.bfloat16(), neg.bfloat16()


def step():
    lat = head.sample_speech_tokens(cond, neg, 1.3, noise=noise)
    return dec.decode(lat.unsqueeze(1), cache=cache, sample_indices=idx, use_cache=True)


def timed(fn, iters=50, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


print(f"n={n} eager: {timed(step) * 1e3:.1f} us per frame-step")
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3):
        step()
torch.cuda.current_stream().wait_stream(s)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
try:
    with torch.cuda.graph(g):
        out = step()
    ref = None
    print(f"n={n} graph: {timed(g.replay) * 1e3:.1f} us per frame-step; output finite: {bool(torch.isfinite(out.float()).all())}")
except Exception as e:  # noqa: BLE001
    print("graph capture failed:", type(e).__name__, str(e)[:300])
