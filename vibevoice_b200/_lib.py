"""ctypes binding of libvvb200.so (include/vvb200.h).

The shared library is the product: if it is missing or fails to load this module raises —
there is no eager / CPU fallback behind it (north_star: "no CPU fallback").
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VVB200_LIB") or os.path.join(_HERE, "libvvb200.so")   # VVB200_LIB: experiment builds

VV_DTYPE_BF16 = 0
VV_DTYPE_F32 = 1
VV_ALGO_DPMPP = 0
VV_ALGO_SDE_DPMPP = 1
VV_DEC_MAX_STAGES = 8

# GemmMode (csrc/vv_gemm.cuh)
GM_STORE_BF16 = 0
GM_GELU_BF16 = 1
GM_SWIGLU_BF16 = 2
GM_RESID_GAMMA_BF16 = 3
GM_PARTIAL_F32 = 4
GM_RESID_GATE_F32 = 5
GM_STORE_F32 = 6
GM_SILU_BF16 = 7



class VVError(RuntimeError):
    pass


class VVHeadDesc(C.Structure):
    _fields_ = [
        ("hidden", C.c_int),
        ("ffn", C.c_int),
        ("latent", C.c_int),
        ("layers", C.c_int),
        ("freq_dim", C.c_int),
        ("eps", C.c_float),
        ("weight_dtype", C.c_int),
        ("max_rows", C.c_int),
    ]


class VVDecDesc(C.Structure):
    _fields_ = [
        ("vae_dim", C.c_int),
        ("n_stages", C.c_int),
        ("channels", C.c_int * VV_DEC_MAX_STAGES),
        ("depths", C.c_int * VV_DEC_MAX_STAGES),
        ("ratios", C.c_int * VV_DEC_MAX_STAGES),
        ("kernel", C.c_int),
        ("eps", C.c_float),
        ("weight_dtype", C.c_int),
        ("max_rows", C.c_int),
    ]


class VVLaunchRec(C.Structure):
    _fields_ = [("M", C.c_int), ("N", C.c_int), ("K", C.c_int), ("bn", C.c_int), ("splits", C.c_int),
                ("mode", C.c_int), ("ms", C.c_float), ("kind", C.c_int), ("bytes", C.c_longlong)]


_vp = C.c_void_p
_i = C.c_int
_f = C.c_float

# name -> (restype, argtypes); must list every symbol include/vvb200.h declares.
SIGNATURES = {
    "vv_version": (_i, []),
    "vv_last_error": (C.c_char_p, []),
    "vv_launch_count": (C.c_longlong, []),
    "vv_host_sync_count": (C.c_longlong, []),
    "vv_timing_enable": (_i, [_i]),
    "vv_timing_fetch": (_i, [C.POINTER(VVLaunchRec), _i]),
    "vv_linear": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _i, _vp, _i, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _i, _i, _vp]),
    "vv_linear_splitk_static": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _i, _vp, _vp, _i, _i, _vp]),
    "vv_linear_plan": (_i, [_i, _i, _i, _i, C.POINTER(_i), C.POINTER(_i), C.POINTER(_i), C.POINTER(_i)]),
    "vv_set_option": (_i, [C.c_char_p, _i]),
    "vv_debug_trace": (_i, [C.POINTER(C.c_longlong), _i]),
    "vv_head_create": (_i, [C.POINTER(VVHeadDesc), C.POINTER(_vp), _i, C.POINTER(_vp)]),
    "vv_head_destroy": (None, [_vp]),
    "vv_head_set_schedule": (_i, [_vp, _i, C.POINTER(_f), C.POINTER(_f), _i, _vp]),
    "vv_head_forward": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _vp]),
    "vv_head_sample": (_i, [_vp, _vp, _vp, _vp, _vp, _f, _vp, _i, _i, _vp]),
    "vv_dec_create": (_i, [C.POINTER(VVDecDesc), C.POINTER(_vp), _i, C.POINTER(_vp)]),
    "vv_dec_destroy": (None, [_vp]),
    "vv_dec_decode": (_i, [_vp, _vp, _i, _i, _i, _f, _f, _vp, _i, _vp]),
    "vv_cache_create": (_i, [_vp, _i, C.POINTER(_vp)]),
    "vv_cache_destroy": (None, [_vp]),
    "vv_cache_zero": (_i, [_vp, _vp, _i, _vp]),
    "vv_cache_zero_all": (_i, [_vp, _vp]),
    "vv_dec_step": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _f, _f, _vp, _i, _vp]),
    "vv_enc_create": (_i, [C.POINTER(VVDecDesc), C.POINTER(_vp), _i, C.POINTER(_vp)]),
    "vv_enc_destroy": (None, [_vp]),
    "vv_enc_encode": (_i, [_vp, _vp, _i, _i, _vp, _i, _vp]),
    "vv_enc_prefill": (_i, [_vp, _vp, _i, C.c_longlong, _vp, _i, _vp]),
    "vv_latent_sample": (_i, [_vp, _vp, _vp, _f, _f, _f, _i, _i, _i, _i, _vp, _vp, _vp]),
    "vv_frame_select": (_i, [_vp, _vp, _i, C.c_int64, C.c_int64, _vp, _vp, _vp, _vp, _vp]),
    "vv_ring_append": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _i, _vp]),
    "vv_embed_scatter": (_i, [_vp, _vp, _i, _i, _vp, _i, _vp]),
    "vv_enc_step": (_i, [_vp, _vp, _vp, _vp, _i, _i, _vp, _i, _vp]),
    "vv_conn_create": (_i, [_i, _i, _f, _i, _i, C.POINTER(_vp), _i, C.POINTER(_vp)]),
    "vv_conn_destroy": (None, [_vp]),
    "vv_conn_forward": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp]),
}

_lib = None


# measurement probes: exported only by a `make PROBE=1` build (include/vvb200_probe.h)
PROBE_SIGNATURES = {
    "vv_debug_mma_commit": (_i, [_i, _i, _i, _i, _vp, _i, _i]),
    "vv_debug_tma_stream": (_i, [_vp, C.c_longlong, _i, _i, _i, _i, _i, _vp]),
}


def load() -> C.CDLL:
    """Load libvvb200.so (built in-tree by __graft_entry__.build()). Raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise VVError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). vibevoice_b200 has no fallback path."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing: fail loudly
        fn.restype = res
        fn.argtypes = args
    for name, (res, args) in PROBE_SIGNATURES.items():
        fn = getattr(lib, name, None)   # present only in a PROBE=1 build
        if fn is not None:
            fn.restype = res
            fn.argtypes = args
    _lib = lib
    return lib


def check(status: int, what: str = "") -> None:
    if status != 0:
        msg = load().vv_last_error().decode("utf-8", "replace")
        raise VVError(f"{what or 'libvvb200'} failed ({status}): {msg}")


def ptr(t) -> C.c_void_p:
    """Device pointer of a torch tensor (or None)."""
    if t is None:
        return C.c_void_p(0)
    return C.c_void_p(t.data_ptr())


def current_stream() -> C.c_void_p:
    import torch

    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
