#!/usr/bin/env python
"""Trace build only: per-phase clock stamps of the persistent head kernel (first 24 phases)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from vibevoice_b200 import _lib, synth  # noqa: E402
from vibevoice_b200.head import DiffusionHeadB200  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
lib = _lib.load()
if len(sys.argv) > 2:
    lib.vv_set_option(b"chain_wstages", int(sys.argv[2]))
if len(sys.argv) > 3:
    lib.vv_set_option(b"chain_stages", int(sys.argv[3]))
cfg = synth.HEAD_1P5B
head = DiffusionHeadB200(synth.init_head(cfg), cfg, max_rows=2 * n)
cond, neg, noise = synth.synth_inputs(n, cfg.hidden, cfg.latent)
cond, neg = cond.cuda().bfloat16(), neg.cuda().bfloat16()
for _ in range(3):
    head.sample_speech_tokens(cond, neg, 1.3, num_steps=10, noise=noise)
torch.cuda.synchronize()
lib.vv_set_option(b"trace_select", 2)
NS = 384
N2 = 160 * 24 * 3 * 16
buf = (C.c_longlong * (160 * NS + N2))()
k = lib.vv_debug_trace(buf, 160 * NS + N2)
names = (["proj"] + sum([["norm", "gu", "down"] for _ in range(4)], []) + ["tail"]
         + ["gu", "down"] + sum([["norm", "gu", "down"] for _ in range(3)], []) + ["tail"])
if k > 0:
    t = np.array(buf[:160 * NS]).reshape(160, NS)[:148].reshape(148, 24, 16)
    def st(v):
        return f"{int(v.min()):6d} {int(np.median(v)):6d} {int(v.max()):6d}"
    print("per phase over CTAs, min/med/max clk:")
    print("  work = done - prev exit | cta-bar = thread 64 waits for its own CTA | poll = flag store -> all seen | "
          "exit = all seen -> thread 64 out | prod/mma = their arrival relative to thread 64's")
    for ph in range(1, 24):
        r = t[:, ph]
        work = r[:, 0] - t[:, ph - 1, 4]
        print(f" {ph:2d} {names[ph]:6s} work {st(work)} | cta-bar {st(r[:, 1] - r[:, 0])} | poll {st(r[:, 3] - r[:, 2])} | "
              f"exit {st(r[:, 4] - r[:, 3])} | prod {st(r[:, 5] - r[:, 0])} | mma {st(r[:, 6] - r[:, 0])}")
    print("inside row phases (thread 64 of CTAs that had work), clk since phase start: median of stamps 8..12")
    for ph in range(1, 24):
        if names[ph] in ("norm", "tail"):
            r = t[:, ph]
            start = t[:, ph - 1, 4]
            sel = (r[:, 0] - start) > 1000
            if names[ph] == "tail":
                sel = np.arange(148) < n            # the CTAs that own a row pair
            print(f" {ph:2d} {names[ph]:5s}", [int(np.median((r[sel, k] - start[sel]))) for k in range(8, 13)], "done", int(np.median(r[sel, 0] - start[sel])))
    print("GEMM phases: MMA warp, clk since phase start (median): W0 ready, A0 ready, last kb ready, issued all, epilogue saw acc_full, tcgen05.ld done | total clk waiting for W, for A | done")
    for ph in range(1, 24):
        if names[ph] in ("gu", "down"):
            r = t[:, ph]
            start = t[:, ph - 1, 4]
            sel = (r[:, 0] - start) > 1000
            print(f" {ph:2d} {names[ph]:5s}", [int(np.median((r[sel, k] - start[sel]))) for k in (13, 8, 11, 12, 14, 15)], "| wait W/A", int(np.median(r[sel, 9])), int(np.median(r[sel, 10])), "| done", int(np.median(r[sel, 0] - start[sel])))
    g = t[:, :, 7]
    print("globaltimer: exit skew over CTAs per phase (ns):", [int(g[:, ph].max() - g[:, ph].min()) for ph in range(24)])
    print("globaltimer: phase length (ns, median CTA):", [int(np.median(g[:, ph] - g[:, ph - 1])) for ph in range(1, 24)])
    print("clock64 span / globaltimer span per CTA (GHz): min/max",
          ((t[:, 23, 4] - t[:, 0, 4]) / (g[:, 23] - g[:, 0])).min(), ((t[:, 23, 4] - t[:, 0, 4]) / (g[:, 23] - g[:, 0])).max())

if k > 160 * NS:
    t2 = np.array(buf[160 * NS:160 * NS + N2]).reshape(160, 24, 3, 16)[:148]
    print("per ring group, clk since the phase started (median over CTAs): weight load issued | weights landed (seen by the MMA warp) | activations landed")
    for ph in (5, 6, 8, 9):
        start = t[:, ph - 1, 4]
        ng = 12 if names[ph] == "gu" else 6
        for what, label in ((2, "W issued"), (0, "W landed"), (1, "A landed")):
            print(f" {ph:2d} {names[ph]:5s} {label}:", [int(np.median(t2[:, ph, what, j] - start)) for j in range(ng)])
    print("epilogue thread 0, clk since the phase started (median): accumulator in registers | gu: first 16 columns computed, second 16 computed, stores issued; down: tile transposed and stored, smem fenced | phase work done")
    for ph in (5, 6, 8, 9):
        start = t[:, ph - 1, 4]
        sel = (t[:, ph, 0] - start) > 1000
        print(f" {ph:2d} {names[ph]:5s}", [int(np.median(t2[sel, ph, 0, j] - start[sel])) for j in (12, 13, 14, 15)], "done", int(np.median(t[sel, ph, 0] - start[sel])))
    ph = 13
    start = t[:, ph - 1, 4]
    sel = np.arange(148) < 2 * n
    print("split step boundary, thread 0, clk since the phase started (median over the CTAs that own half a pair): 0 state read | 1 rows issued | 3 folded + summed | 4 (1) posted | 5 (1) received | 6 modulated | 7 first weight slot | 8 final Linear loop | 9 partials summed | 10 (2) received | 11 solver done | 13 input projection done | 12 norm sum | 14 (3) received | 15 end")
    print("  ", [(j, int(np.median(t2[sel, ph, 1, j] - start[sel]))) for j in (0, 1, 3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 12, 14, 15)])
    print("   inside the two rowc_issue calls (thread 0): before expect_tx, after it, after its partial copy:", [int(np.median(t2[sel, ph, 2, j] - start[sel])) for j in (0, 1, 2, 4, 5, 6)])
