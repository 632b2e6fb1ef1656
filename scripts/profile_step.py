#!/usr/bin/env python
"""One frame-step of bench.py's workload inside a cudaProfilerStart/Stop window.

    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
        --log-file gpurun_out/launches.csv python scripts/profile_step.py --steps 1

Without ncu it prints the CUDA-event time of the same window (the numbers ncu prints are
cold-cache and serialised; compare shares, not absolutes).
"""
from __future__ import annotations

import argparse
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from vibevoice_b200 import _lib, synth  # noqa: E402
from vibevoice_b200._lib import check, ptr  # noqa: E402
from vibevoice_b200.decoder import AcousticDecoderB200  # noqa: E402
from vibevoice_b200.head import DiffusionHeadB200  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=1)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--part", default="both", choices=["both", "head", "dec", "tail", "all"])
    ap.add_argument("--ddpm-steps", type=int, default=10)
    ap.add_argument("--opt", action="append", default=[], help="name=value for vv_set_option (repeatable)")
    ap.add_argument("--log", action="store_true", help="print kernel name + warm device time of every vv_launch of one step (stderr)")
    ap.add_argument("--launches", action="store_true", help="print the per-launch GEMM timing records of one step")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    n = args.batch
    hcfg, dcfg = synth.HEAD_1P5B, synth.DEC_FULL
    lib = _lib.load()
    for o in args.opt:
        k, v = o.split("=")
        lib.vv_set_option(k.encode(), int(v))
    head = DiffusionHeadB200(synth.init_head(hcfg), hcfg, device=dev, max_rows=2 * n)
    dec = AcousticDecoderB200(synth.init_decoder(dcfg), dcfg, device=dev, max_rows=n)
    head.set_schedule(args.ddpm_steps, "dpmsolver++")
    cache = dec.new_cache(n)
    idx = torch.arange(n, device=dev)
    stp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    cond, neg, noise = synth.synth_inputs(n, hcfg.hidden, hcfg.latent)
    cond, neg, noise = cond.to(dev).bfloat16(), neg.to(dev).bfloat16(), noise.to(dev)
    lat = torch.randn(n, hcfg.latent, device=dev).bfloat16()
    wave = torch.empty(n, 1, dcfg.hop, device=dev, dtype=torch.bfloat16)
    if args.part in ("tail", "all"):      # rest of the frame iteration: semantic encode + both connectors
        from vibevoice_b200.encoder import SemanticEncoderB200, SpeechConnectorB200
        ecfg = synth.ENC_FULL
        enc = SemanticEncoderB200(synth.init_encoder(ecfg), ecfg, device=dev, max_rows=n)
        ecache = enc.new_cache(n)
        ca, cs = synth.ConnCfg(64, hcfg.hidden), synth.ConnCfg(ecfg.dim, hcfg.hidden)
        conn_a = SpeechConnectorB200(synth.init_connector(ca, tag="a"), ca, device=dev, max_rows=n)
        conn_s = SpeechConnectorB200(synth.init_connector(cs, tag="s"), cs, device=dev, max_rows=n)
        feat = torch.empty(n, 1, ecfg.dim, device=dev, dtype=torch.bfloat16)
        emb_a = torch.empty(n, hcfg.hidden, device=dev, dtype=torch.bfloat16)
        emb = torch.empty(n, hcfg.hidden, device=dev, dtype=torch.bfloat16)
        wave.normal_()

    def step():
        if args.part in ("both", "head", "all"):
            check(lib.vv_head_sample(head._h, ptr(cond), ptr(neg), ptr(noise), None, 1.3, ptr(lat), n,
                                     _lib.VV_DTYPE_BF16, stp), "vv_head_sample")
        if args.part in ("both", "dec", "all"):
            check(lib.vv_dec_step(dec._h, cache._c, ptr(lat), ptr(idx), n, 1, 1, 1.0, 0.0, ptr(wave),
                                  _lib.VV_DTYPE_BF16, stp), "vv_dec_step")

        if args.part in ("tail", "all"):
            check(lib.vv_enc_step(enc._h, ecache._c, ptr(wave), ptr(idx), n, 1, ptr(feat), _lib.VV_DTYPE_BF16, stp), "vv_enc_step")
            check(lib.vv_conn_forward(conn_a._h, ptr(lat), n, _lib.VV_DTYPE_BF16, None, ptr(emb_a), stp), "vv_conn_forward")
            check(lib.vv_conn_forward(conn_s._h, ptr(feat), n, _lib.VV_DTYPE_BF16, ptr(emb_a), ptr(emb), stp), "vv_conn_forward")

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    rt = torch.cuda.cudart()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    rt.cudaProfilerStart()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    rt.cudaProfilerStop()
    print(f"part={args.part} batch={n} steps={args.steps} opts={args.opt}: {e0.elapsed_time(e1) / args.steps:.4f} ms/step")
    if args.log:
        lib.vv_set_option(b"launch_log", 1)
        step()
        lib.vv_set_option(b"launch_log", 0)      # prints one line per launch to stderr
    if args.launches:
        lib.vv_timing_enable(1)
        step()
        torch.cuda.synchronize()
        recs = (_lib.VVLaunchRec * 4096)()
        k = lib.vv_timing_fetch(recs, 4096)
        lib.vv_timing_enable(0)
        for r in recs[:k]:
            print(f"  kind={r.kind} M={r.M} N={r.N} K={r.K} bn={r.bn} sp={r.splits} mode={r.mode}: {r.ms * 1e3:.2f} us")


if __name__ == "__main__":
    main()
