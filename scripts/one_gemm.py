#!/usr/bin/env python
"""One GEMM shape with a fixed launch plan, a few launches over rotating weight copies — the
target of `ncu --set full` captures (profiles/)."""
import argparse
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from vibevoice_b200 import _lib  # noqa: E402
from vibevoice_b200._lib import ptr  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--M", type=int, default=128)
ap.add_argument("--N", type=int, default=9216)
ap.add_argument("--K", type=int, default=1536)
ap.add_argument("--mode", type=int, default=2)
ap.add_argument("--bn", type=int, default=0)
ap.add_argument("--splits", type=int, default=0)
ap.add_argument("--cluster", type=int, default=0)
ap.add_argument("--stages", type=int, default=0)
ap.add_argument("--iters", type=int, default=12)
ap.add_argument("--persistent", type=int, default=1)
a = ap.parse_args()
lib = _lib.load()
lib.vv_set_option(b"gemm_persistent", a.persistent)
dev = torch.device("cuda:0")
M, N, K = a.M, a.N, a.K
copies = int(300e6 // (N * K * 2)) + 1
x = torch.randn(M, K, device=dev).bfloat16()
ws = [torch.randn(N, K, device=dev).bfloat16() / K ** 0.5 for _ in range(copies)]
bn, sp, cl, st = C.c_int(a.bn), C.c_int(a.splits), C.c_int(a.cluster), C.c_int(a.stages)
lib.vv_linear_plan(M, N, K, a.mode, C.byref(bn), C.byref(sp), C.byref(cl), C.byref(st))
if a.mode == 4:
    out = torch.zeros(sp.value, M, N, device=dev)
    ldo = N
elif a.mode == 2:
    out = torch.zeros(M, N // 2, device=dev, dtype=torch.bfloat16)
    ldo = N // 2
else:
    out = torch.zeros(M, N, device=dev, dtype=torch.bfloat16)
    ldo = N
stream = torch.cuda.Stream()
stp = C.c_void_p(stream.cuda_stream)
with torch.cuda.stream(stream):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for it in range(2):
        e0.record(stream)
        for i in range(a.iters):
            rc = lib.vv_linear(ptr(x), K, ptr(ws[i % copies]), K, M, N, K, a.mode, ptr(out), ldo, None, None, None, 0,
                               None, 0, bn.value, sp.value, cl.value, st.value, stp)
            assert rc == 0, lib.vv_last_error()
        e1.record(stream)
        stream.synchronize()
print(f"persistent={a.persistent} M={M} N={N} K={K} mode={a.mode} bn={bn.value} sp={sp.value} cl={cl.value} st={st.value}: "
      f"{e0.elapsed_time(e1) * 1e3 / a.iters:.2f} us/launch")
buf = (C.c_longlong * (160 * 48))()
n = lib.vv_debug_trace(buf, 160 * 48)
if n > 0:
    import numpy as np
    t = np.array(buf[:n]).reshape(160, 48)
    kb = (K + 63) // 64 // max(sp.value, 1)
    for cta in (0, 1, 70, 143):
        r = t[cta]
        base = r[0]
        print(f"cta {cta}: setup_done {r[1]-base}  B-prefetch issued {r[2]-base}  griddep_wait done {r[3]-base}  "
              f"accum ready {r[4]-base}  epilogue done {r[5]-base}")
        print("   full-barrier times:", [int(r[8 + i] - base) for i in range(min(kb, 40))])
