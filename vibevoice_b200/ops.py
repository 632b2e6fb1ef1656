"""Thin operator wrappers over the C ABI (tensor in / tensor out)."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import check, current_stream, ptr


def linear(a, w, mode=_lib.GM_STORE_BF16, bias=None, gamma=None, resid=None, gate=None,
           out=None, bn=0, splits=0, cluster=0, stages=0):
    """``epilogue(a @ w.T)`` on the tcgen05 GEMM (vv_linear).

    a [M,K] bf16, w [N,K] bf16 (nn.Linear layout: modular_vibevoice_diffusion_head.py:59-62).
    Returns the output tensor; for GM_PARTIAL_F32 returns [splits, M, N] fp32 partial sums.
    """
    lib = _lib.load()
    assert a.is_cuda and w.is_cuda and a.dtype == torch.bfloat16 and w.dtype == torch.bfloat16
    assert a.stride(-1) == 1 and w.stride(-1) == 1
    M, K = a.shape
    N = w.shape[0]
    bn_c, sp_c, cl_c, st_c = C.c_int(bn), C.c_int(splits), C.c_int(cluster), C.c_int(stages)
    check(lib.vv_linear_plan(M, N, K, mode, C.byref(bn_c), C.byref(sp_c), C.byref(cl_c), C.byref(st_c)),
          "vv_linear_plan")
    if out is None:
        if mode == _lib.GM_PARTIAL_F32:
            out = torch.empty(sp_c.value, M, N, device=a.device, dtype=torch.float32)
        elif mode in (_lib.GM_RESID_GATE_F32, _lib.GM_STORE_F32):
            out = torch.empty(M, N, device=a.device, dtype=torch.float32)
        elif mode == _lib.GM_SWIGLU_BF16:
            out = torch.empty(M, N // 2, device=a.device, dtype=torch.bfloat16)
        else:
            out = torch.empty(M, N, device=a.device, dtype=torch.bfloat16)
    ldo = out.stride(-2)
    check(
        lib.vv_linear(ptr(a), a.stride(0), ptr(w), w.stride(0), M, N, K, mode, ptr(out), ldo,
                      ptr(bias), ptr(gamma), ptr(resid), resid.stride(0) if resid is not None else 0,
                      ptr(gate), gate.stride(0) if gate is not None else 0, bn_c.value, sp_c.value,
                      cl_c.value, st_c.value, current_stream()),
        "vv_linear",
    )
    return out


SPLITK_STATIC_MAX_FLOATS = 1 << 20   # include/vvb200.h: VV_SPLITK_STATIC_MAX_FLOATS


def linear_splitk_static(a, w, out, bias=None, splits=0):
    """``out = a @ w.T + bias`` through vv_linear_splitk_static: split-K partials in the library's static device
    array, no allocation anywhere in the call (``out`` is the caller's, bf16 or fp32 [M,N]).  No counterpart in the
    reference (added for callers that cannot call cudaMalloc); splits*M*N is limited to SPLITK_STATIC_MAX_FLOATS."""
    lib = _lib.load()
    assert a.is_cuda and w.is_cuda and out.is_cuda and a.dtype == torch.bfloat16 and w.dtype == torch.bfloat16
    assert a.stride(-1) == 1 and w.stride(-1) == 1 and out.stride(-1) == 1
    M, K = a.shape
    N = w.shape[0]
    assert out.shape == (M, N) and out.dtype in (torch.bfloat16, torch.float32)
    dt = _lib.VV_DTYPE_BF16 if out.dtype == torch.bfloat16 else _lib.VV_DTYPE_F32
    check(lib.vv_linear_splitk_static(ptr(a), a.stride(0), ptr(w), w.stride(0), M, N, K, splits, ptr(bias), ptr(out),
                                      out.stride(0), dt, current_stream()), "vv_linear_splitk_static")
    return out
