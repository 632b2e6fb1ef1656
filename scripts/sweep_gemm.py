#!/usr/bin/env python
"""Sweep of the tcgen05 GEMM's launch plan (tile width, split-K, cluster multicast, ring depth, PDL)
on the weight-streaming shapes of the per-frame path.  Weights rotate over enough copies to exceed
the 126 MB L2, launches are back to back on one stream (what a frame-step looks like), time is CUDA
events over the whole loop.  Prints GB/s of weight bytes and checks each config against torch."""
from __future__ import annotations

import argparse
import ctypes as C
import itertools
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from vibevoice_b200 import _lib  # noqa: E402
from vibevoice_b200._lib import ptr  # noqa: E402

SHAPES = {
    # name: (M, N, K, mode)
    "head_gu": (128, 9216, 1536, _lib.GM_SWIGLU_BF16),
    "head_down": (128, 1536, 4608, _lib.GM_PARTIAL_F32),
    "head_final": (128, 64, 1536, _lib.GM_PARTIAL_F32),
    "dec0_ffn1": (64, 8192, 2048, _lib.GM_GELU_BF16),
    "dec0_ffn2": (64, 2048, 8192, _lib.GM_PARTIAL_F32),
    "dec1_up": (64, 8192, 4096, _lib.GM_STORE_BF16),
    "dec1_ffn1": (512, 4096, 1024, _lib.GM_GELU_BF16),
    "dec1_ffn2": (512, 1024, 4096, _lib.GM_STORE_BF16),
    "dec2_ffn1": (2560, 2048, 512, _lib.GM_GELU_BF16),
    "dec2_ffn2": (2560, 512, 2048, _lib.GM_STORE_BF16),
    "ada_all": (1280, 21504, 1536, _lib.GM_STORE_BF16),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="head_gu,head_down,dec0_ffn1,dec0_ffn2")
    ap.add_argument("--iters", type=int, default=40)
    ap.add_argument("--stream", default="new", choices=["new", "default"])
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    lib = _lib.load()
    dev = torch.device("cuda:0")
    stream = torch.cuda.Stream() if args.stream == "new" else torch.cuda.current_stream()
    stp = C.c_void_p(stream.cuda_stream)
    for name in args.shapes.split(","):
        M, N, K, mode = SHAPES[name]
        wbytes = N * K * 2
        copies = max(2, int(300e6 // wbytes) + 1)
        g = torch.Generator().manual_seed(0)
        a = (torch.randn(M, K, generator=g) * 0.5).bfloat16().to(dev)
        ws = [(torch.randn(N, K, generator=g) / K ** 0.5).bfloat16().to(dev) for _ in range(min(copies, 2))]
        while len(ws) < copies:
            ws.append(ws[len(ws) % 2].clone())
        ref = [a.float() @ w.float().t() for w in ws[:2]]
        print(f"== {name}: M={M} N={N} K={K} mode={mode}  weights {wbytes / 1e6:.1f} MB x {copies} copies", flush=True)
        bns = [32, 64, 128, 256] if not args.quick else [64, 128]
        cls = [1, 2, 4]
        sts = [4, 8] if not args.quick else [4]
        splits_opts = [0] if mode != _lib.GM_PARTIAL_F32 else [0, 2, 3, 4, 6, 8, 12]
        results = []
        for bn, cl, st, sp, pdl in itertools.product(bns, cls, sts, splits_opts, [1, 0]):
            if bn > N:
                continue
            bn_c, sp_c, cl_c, st_c = C.c_int(bn), C.c_int(sp), C.c_int(cl), C.c_int(st)
            lib.vv_linear_plan(M, N, K, mode, C.byref(bn_c), C.byref(sp_c), C.byref(cl_c), C.byref(st_c))
            splits = sp_c.value
            kb = (K + 63) // 64
            if splits > 1 and (kb + splits - 1) // splits * (splits - 1) >= kb:
                continue
            n_ctas = -(-(-(-N // bn)) // cl) * cl * -(-M // 128) * splits
            if n_ctas < 64 or n_ctas > 148 * 6:
                continue
            with torch.cuda.stream(stream):     # the zero-fill must be ordered before the GEMM
                if mode == _lib.GM_PARTIAL_F32:
                    out = torch.zeros(splits, M, N, device=dev)
                    ldo = N
                elif mode == _lib.GM_SWIGLU_BF16:
                    out = torch.zeros(M, N // 2, device=dev, dtype=torch.bfloat16)
                    ldo = N // 2
                else:
                    out = torch.zeros(M, N, device=dev, dtype=torch.bfloat16)
                    ldo = N
            lib.vv_set_option(b"pdl", pdl)

            def call(w):
                rc = lib.vv_linear(ptr(a), K, ptr(w), K, M, N, K, mode, ptr(out), ldo, None, None, None, 0, None, 0,
                                   bn, splits, cl, st_c.value, stp)
                if rc != 0:
                    raise RuntimeError(lib.vv_last_error().decode())

            with torch.cuda.stream(stream):
                try:
                    call(ws[1])
                    stream.synchronize()
                except RuntimeError as e:
                    print(f"   bn={bn} cl={cl} st={st} sp={splits}: ERROR {e}")
                    continue
                # correctness
                if mode == _lib.GM_PARTIAL_F32:
                    got = out.sum(0)
                    want = ref[1]
                elif mode == _lib.GM_SWIGLU_BF16:
                    got = out.float()
                    want = torch.nn.functional.silu(ref[1][:, 0::2]) * ref[1][:, 1::2]
                elif mode == _lib.GM_GELU_BF16:
                    got = out.float()
                    want = torch.nn.functional.gelu(ref[1])
                else:
                    got, want = out.float(), ref[1]
                err = (got - want).abs().max().item() / max(1.0, want.abs().max().item())
                ok = err < 1e-2
                for i in range(5):
                    call(ws[i % copies])
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for i in range(args.iters):
                    call(ws[i % copies])
                e1.record(stream)
                stream.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / args.iters
            results.append((us, bn, cl, st_c.value, splits, pdl, n_ctas, ok, err))
        lib.vv_set_option(b"pdl", 1)
        results.sort()
        for us, bn, cl, st, sp, pdl, n_ctas, ok, err in results[:14]:
            print(f"   {us:7.2f} us  {wbytes / us / 1e3:7.1f} GB/s  bn={bn:3d} cl={cl} st={st} sp={sp:2d} pdl={pdl} ctas={n_ctas:4d} "
                  f"{'ok' if ok else 'MISMATCH %.3g' % err}")
        bad = [r for r in results if not r[7]]
        if bad:
            print(f"   !! {len(bad)} configs MISMATCH, e.g. {bad[0]}")
        # default plan
        bn_c, sp_c, cl_c, st_c = C.c_int(0), C.c_int(0), C.c_int(0), C.c_int(0)
        lib.vv_linear_plan(M, N, K, mode, C.byref(bn_c), C.byref(sp_c), C.byref(cl_c), C.byref(st_c))
        mine = [r for r in results if (r[1], r[2], r[3], r[4], r[5]) == (bn_c.value, cl_c.value, st_c.value, sp_c.value, 1)]
        print(f"   planner default: bn={bn_c.value} cl={cl_c.value} st={st_c.value} sp={sp_c.value}"
              + (f" -> {mine[0][0]:.2f} us" if mine else " (not in sweep)"), flush=True)


if __name__ == "__main__":
    main()
