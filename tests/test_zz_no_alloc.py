"""vv_linear_splitk_static (no counterpart in the reference; added for callers that cannot call cudaMalloc): the
split-K linear whose partial sums live in the library's static device array, against the existing entry point
vv_linear on the same inputs.  Collected last on purpose (pytest -x: a failure here hides nothing after it).

Tolerance: the one tests/test_gemm_gpu.py uses for the same GEMM — bf16 inputs on both sides, so only the fp32
summation order (and the final bf16 rounding) differ."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from vibevoice_b200 import _lib, ops  # noqa: E402
from vibevoice_b200._lib import VVError  # noqa: E402


def _mk(M, N, K, seed):
    g = torch.Generator().manual_seed(seed)
    a = (torch.randn(M, K, generator=g) * 0.5).bfloat16().cuda()
    w = (torch.randn(N, K, generator=g) * (1.0 / K ** 0.5)).bfloat16().cuda()
    bias = (torch.randn(N, generator=g) * 0.1).cuda()
    return a, w, bias


@pytest.mark.parametrize("M,N,K,splits", [(64, 2048, 8192, 8), (2, 64, 1536, 0), (128, 1536, 4608, 4), (77, 96, 72, 0),
                                          (64, 128, 14336, 32)])
def test_static_splitk_matches_vv_linear_two_calls_in_a_row(M, N, K, splits):
    assert max(splits, 1) * M * N <= ops.SPLITK_STATIC_MAX_FLOATS
    out32 = torch.empty(M, N, device="cuda", dtype=torch.float32)
    out16 = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    for call in range(2):                      # the second call finds the first call's partials in the array
        a, w, bias = _mk(M, N, K, seed=10 + call)
        ref32 = ops.linear(a, w, mode=_lib.GM_STORE_F32, bias=bias)         # existing entry point, one launch
        ref16 = ops.linear(a, w, mode=_lib.GM_STORE_BF16, bias=bias)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        ops.linear_splitk_static(a, w, out32, bias=bias, splits=splits)
        ops.linear_splitk_static(a, w, out16, bias=bias, splits=splits)
        torch.cuda.synchronize()
        if call == 1:                          # steady state: the call took no device memory
            assert torch.cuda.mem_get_info()[0] == free0
        scale = max(1.0, ref32.abs().max().item())
        err = (out32 - ref32).abs().max().item()
        assert err < 2e-3 * scale, f"call {call}: fp32 max err {err}"
        tol = 2e-3 + 8e-3 * ref32.abs()
        assert ((out16.float() - ref32).abs() <= tol).all(), f"call {call}: bf16 {(out16.float() - ref32).abs().max()}"
        assert ((out16.float() - ref16.float()).abs() <= tol).all()
    # no bias, strided output rows
    a, w, _ = _mk(M, N, K, seed=12)
    wide = torch.full((M, N + 8), 7.0, device="cuda", dtype=torch.float32)
    ops.linear_splitk_static(a, w, wide[:, :N], splits=splits)
    ref = ops.linear(a, w, mode=_lib.GM_STORE_F32)
    torch.cuda.synchronize()
    assert (wide[:, :N] - ref).abs().max().item() < 2e-3 * max(1.0, ref.abs().max().item())
    assert (wide[:, N:] == 7.0).all()


def test_static_splitk_rejects_what_does_not_fit():
    a, w, bias = _mk(256, 4096, 512, seed=3)
    out = torch.empty(256, 4096, device="cuda", dtype=torch.float32)
    with pytest.raises(VVError, match="exceeds the static scratch"):
        ops.linear_splitk_static(a, w, out, bias=bias, splits=2)          # 2 * 256 * 4096 = 2 Mi floats > 1 Mi
    ops.linear_splitk_static(a, w, out, bias=bias, splits=0)              # chosen by the library: shrunk to fit
    ref = ops.linear(a, w, mode=_lib.GM_STORE_F32, bias=bias)
    torch.cuda.synchronize()
    assert (out - ref).abs().max().item() < 2e-3 * max(1.0, ref.abs().max().item())
    big = torch.empty(2048, 4096, device="cuda", dtype=torch.float32)
    a2, w2, _ = _mk(2048, 4096, 64, seed=4)
    with pytest.raises(VVError, match="exceeds the static scratch"):
        ops.linear_splitk_static(a2, w2, big)                             # does not fit even unsplit
    lib = _lib.load()
    try:
        lib.vv_timing_enable(1)
        with pytest.raises(VVError, match="vv_timing_enable"):
            ops.linear_splitk_static(a, w, out, bias=bias)
    finally:
        lib.vv_timing_enable(0)
