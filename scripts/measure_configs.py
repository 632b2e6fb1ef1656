#!/usr/bin/env python
"""BASELINE.json configs other than the bench.py headline (configs[1], streaming): one table.

  2(O)  1.5B, 64 utterances x 450 frames with every condition known up front (offline): head on
        M = 2*28800 rows, decoder [64, 450, 64] time-tiled inside the library
  4     VibeVoice-Large head (H = 3584), n = 1, per frame; decoder over 20 250 frames (45 min), offline
  5     n = 1 streaming frame-step latency, head and decoder separately

CUDA events on the current stream, warm-up first, synthetic inputs, random-init weights (synth.py)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from vibevoice_b200 import _lib, synth  # noqa: E402
from vibevoice_b200.decoder import AcousticDecoderB200  # noqa: E402
from vibevoice_b200.head import DiffusionHeadB200  # noqa: E402


def timed(fn, iters, warmup):
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--skip", default="", help="comma list of configs to skip: 2o,4,5")
    ap.add_argument("--opt", action="append", default=[], help="name=value for vv_set_option (repeatable)")
    args = ap.parse_args()
    for o in args.opt:
        k, v = o.split("=")
        _lib.load().vv_set_option(k.encode(), int(v))
    print(json.dumps(measure(torch.device("cuda:0"), set(args.skip.split(","))), indent=1))


def measure(dev, skip=frozenset()):
    """The three measurements as a dict (bench.py emits it under `other_configs`, so they carry its clock record)."""
    out = {}
    dcfg = synth.DEC_FULL
    Wd = synth.init_decoder(dcfg, seed=1234)
    FR = 7.5

    if "5" not in skip:
        hcfg = synth.HEAD_1P5B
        head = DiffusionHeadB200(synth.init_head(hcfg, seed=1234), hcfg, device=dev, max_rows=2)
        dec = AcousticDecoderB200(Wd, dcfg, device=dev, max_rows=1)
        head.set_schedule(10, "dpmsolver++")
        cache = dec.new_cache(1)
        idx = torch.zeros(1, dtype=torch.int64, device=dev)
        cond, neg, noise = [t.to(dev) for t in synth.synth_inputs(1, hcfg.hidden, hcfg.latent)]
        cond, neg = cond.bfloat16(), neg.bfloat16()
        lat = torch.randn(1, 1, 64, device=dev).bfloat16()
        t_head = timed(lambda: head.sample_speech_tokens(cond, neg, 1.3, noise=noise), 50, 5)
        t_dec = timed(lambda: dec.decode(lat, cache=cache, sample_indices=idx, use_cache=True), 50, 5)
        out["config5_n1_streaming"] = {"head_us_per_frame": round(t_head * 1e3, 1), "decoder_us_per_frame": round(t_dec * 1e3, 1),
                                       "frame_step_us": round((t_head + t_dec) * 1e3, 1),
                                       "realtime_factor": round((1 / FR) / ((t_head + t_dec) / 1e3), 1),
                                       "hbm_floor_us": 381, "note": "no CUDA graph; python + ctypes call overhead included"}
        del head, dec, cache

    if "4" not in skip:
        hcfg = synth.HEAD_LARGE
        head = DiffusionHeadB200(synth.init_head(hcfg, seed=7), hcfg, device=dev, max_rows=2)
        head.set_schedule(10, "dpmsolver++")
        cond, neg, noise = [t.to(dev) for t in synth.synth_inputs(1, hcfg.hidden, hcfg.latent)]
        cond, neg = cond.bfloat16(), neg.bfloat16()
        t_head = timed(lambda: head.sample_speech_tokens(cond, neg, 1.3, noise=noise), 30, 5)
        del head
        F = 20250
        dec = AcousticDecoderB200(Wd, dcfg, device=dev, max_rows=64)     # 64 frames per kernel pass, tiled in time
        lat = torch.randn(1, F, 64, device=dev).bfloat16()
        t_dec = timed(lambda: dec.decode(lat), 2, 1)
        out["config4_large_45min"] = {"large_head_us_per_frame_n1": round(t_head * 1e3, 1),
                                      "head_total_s_for_20250_frames": round(t_head * F / 1e3, 2),
                                      "decoder_offline_s": round(t_dec / 1e3, 3),
                                      "decoder_audio_s_per_s": round((F / FR) / (t_dec / 1e3), 0),
                                      "decoder_tflops": round(F * 3.632e9 / (t_dec / 1e3) / 1e12, 1)}
        del dec, lat

    if "2o" not in skip:
        hcfg = synth.HEAD_1P5B
        n, F = 64, 450
        rows = n * F
        head = DiffusionHeadB200(synth.init_head(hcfg, seed=1234), hcfg, device=dev, max_rows=2 * rows)
        head.set_schedule(10, "dpmsolver++")
        g = torch.Generator().manual_seed(5)
        cond = torch.randn(rows, hcfg.hidden, generator=g).bfloat16().to(dev)
        neg = torch.randn(rows, hcfg.hidden, generator=g).bfloat16().to(dev)
        noise = torch.randn(2 * rows, hcfg.latent, generator=g).to(dev)
        t_head = timed(lambda: head.sample_speech_tokens(cond, neg, 1.3, noise=noise), 2, 1)
        del head, cond, neg, noise
        torch.cuda.empty_cache()
        dec = AcousticDecoderB200(Wd, dcfg, device=dev, max_rows=64 * 8)  # 8 frames of all 64 utterances per pass
        lat = torch.randn(n, F, 64, device=dev).bfloat16()
        t_dec = timed(lambda: dec.decode(lat), 2, 1)
        audio_s = n * F / FR
        out["config2_offline_64x60s"] = {"head_s": round(t_head / 1e3, 3), "head_tflops": round(rows * 4.736e9 / (t_head / 1e3) / 1e12, 1),
                                         "decoder_s": round(t_dec / 1e3, 3),
                                         "decoder_tflops": round(rows * 3.632e9 / (t_dec / 1e3) / 1e12, 1),
                                         "audio_s_per_s": round(audio_s / ((t_head + t_dec) / 1e3), 0)}
        del dec, lat
    torch.cuda.empty_cache()
    return out


if __name__ == "__main__":
    main()
