"""tcgen05 GEMM (vv_linear) vs. a plain PyTorch fp32 reference of the same op, every epilogue.

Inputs are bf16 on both sides, so the only differences are fp32 accumulation order and the
final bf16 rounding: tolerances are tight on purpose (a wrong swizzle / descriptor shows up
as O(1) error, not 1e-2).
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

from vibevoice_b200 import _lib, ops  # noqa: E402


def _mk(M, N, K, seed=0):
    g = torch.Generator().manual_seed(seed)
    a = (torch.randn(M, K, generator=g) * 0.5).bfloat16().cuda()
    w = (torch.randn(N, K, generator=g) * (1.0 / K ** 0.5)).bfloat16().cuda()
    return a, w


def _ref(a, w):
    return a.float() @ w.float().t()


SHAPES = [
    (128, 256, 64), (128, 128, 128), (2, 64, 1536), (64, 8192, 2048), (128, 9216, 1536),
    (128, 1536, 4608), (300, 512, 448), (1000, 128, 32), (4096, 32, 128), (256, 2560, 2048),
    (130, 1280, 1024), (77, 96, 72), (2560, 512, 2048), (2560, 2048, 512),
]


@pytest.mark.parametrize("M,N,K", SHAPES)
@pytest.mark.parametrize("bn", [0, 32, 64, 128, 256])
def test_store_f32(M, N, K, bn):
    a, w = _mk(M, N, K)
    out = ops.linear(a, w, mode=_lib.GM_STORE_F32, bn=bn)
    torch.cuda.synchronize()
    ref = _ref(a, w)
    err = (out - ref).abs().max().item()
    assert err < 2e-3 * max(1.0, ref.abs().max().item()), f"max err {err}"


@pytest.mark.parametrize("M,N,K", SHAPES)
def test_bias_bf16_gelu_silu(M, N, K):
    a, w = _mk(M, N, K, seed=1)
    bias = torch.randn(N, device="cuda") * 0.1
    ref = _ref(a, w) + bias
    for mode, fn in [(_lib.GM_STORE_BF16, lambda x: x),
                     (_lib.GM_GELU_BF16, torch.nn.functional.gelu),
                     (_lib.GM_SILU_BF16, torch.nn.functional.silu)]:
        out = ops.linear(a, w, mode=mode, bias=bias)
        torch.cuda.synchronize()
        r = fn(ref)
        tol = 2e-3 + 8e-3 * r.abs()
        assert ((out.float() - r).abs() <= tol).all(), f"mode {mode}: {(out.float()-r).abs().max()}"


@pytest.mark.parametrize("M,N,K", [(128, 9216, 1536), (64, 8192, 2048), (2, 64, 1536), (300, 512, 448), (2560, 512, 2048)])
@pytest.mark.parametrize("cluster", [1, 2, 4])
@pytest.mark.parametrize("stages", [1, 3, 8])
def test_cluster_multicast_and_ring_depth(M, N, K, cluster, stages):
    """A-tile multicast across 1/2/4-CTA clusters (grid padded to whole clusters) and every ring depth
    give the same result as the plain launch."""
    a, w = _mk(M, N, K, seed=3)
    ref = _ref(a, w)
    for bn in (32, 128):
        out = ops.linear(a, w, mode=_lib.GM_STORE_F32, bn=bn, cluster=cluster, stages=stages)
        torch.cuda.synchronize()
        err = (out - ref).abs().max().item()
        assert err < 2e-3 * max(1.0, ref.abs().max().item()), f"bn {bn}: max err {err}"


def test_pdl_switch_and_back_to_back_launches():
    """Programmatic dependent launch: a chain of dependent GEMMs (each reads the previous output) is
    correct with PDL on and off."""
    lib = _lib.load()
    a, w = _mk(128, 1536, 1536, seed=4)
    w2 = (w * 0.5).contiguous()
    try:
        outs = []
        for pdl in (1, 0):
            lib.vv_set_option(b"pdl", pdl)
            x = a
            for i in range(6):
                x = ops.linear(x, w if i % 2 == 0 else w2, mode=_lib.GM_STORE_BF16)
            torch.cuda.synchronize()
            outs.append(x.float())
    finally:
        lib.vv_set_option(b"pdl", 1)
    assert torch.equal(outs[0], outs[1])
    ref = a
    for i in range(6):
        ref = (ref.float() @ (w if i % 2 == 0 else w2).float().t()).bfloat16()
    assert (outs[0] - ref.float()).abs().max() <= 2e-2 * max(1.0, ref.float().abs().max().item())


@pytest.mark.parametrize("M,N,K", [(128, 9216, 1536), (2, 9216, 1536), (500, 256, 128)])
def test_swiglu(M, N, K):
    a, w = _mk(M, N, K, seed=2)
    out = ops.linear(a, w, mode=_lib.GM_SWIGLU_BF16)
    torch.cuda.synchronize()
    acc = _ref(a, w)
    r = torch.nn.functional.silu(acc[:, 0::2]) * acc[:, 1::2]
    assert out.shape == (M, N // 2)
    assert ((out.float() - r).abs() <= 2e-3 + 8e-3 * r.abs()).all()


@pytest.mark.parametrize("M,N,K", [(64, 2048, 8192), (128, 1536, 4608), (2, 64, 1536), (700, 128, 512)])
def test_partial_splitk(M, N, K):
    a, w = _mk(M, N, K, seed=3)
    for splits in (0, 1, 3, 6):
        out = ops.linear(a, w, mode=_lib.GM_PARTIAL_F32, splits=splits)
        torch.cuda.synchronize()
        ref = _ref(a, w)
        got = out.sum(0)
        err = (got - ref).abs().max().item()
        assert err < 2e-3 * max(1.0, ref.abs().max().item()), f"splits={splits} -> {out.shape[0]}: {err}"


@pytest.mark.parametrize("M,N,K", [(64, 2048, 8192), (1, 128, 72), (7, 256, 1536), (33, 2048, 16384), (32, 8192, 2048),
                                   (64, 128, 14336), (50, 1024, 200)])
def test_swapped_operands_splitk(M, N, K):
    """M <= 64 rows, split-K partials, vv_set_option("gemm_swap", 1): the planner puts the WEIGHTS on the MMA's 128-row
    side (out^T in TMEM).  Off by default (no measured gain); kept correct.  Same
    sums as the plain tiling (option off) and as the fp32 reference, for ragged M / K and every split count."""
    lib = _lib.load()
    a, w = _mk(M, N, K, seed=5)
    ref = _ref(a, w)
    tol = 2e-3 * max(1.0, ref.abs().max().item())
    try:
        for splits in (0, 1, 2, 5):
            lib.vv_set_option(b"gemm_swap", 1)
            sw = ops.linear(a, w, mode=_lib.GM_PARTIAL_F32, splits=splits)
            lib.vv_set_option(b"gemm_swap", 0)
            pl = ops.linear(a, w, mode=_lib.GM_PARTIAL_F32, splits=splits)
            torch.cuda.synchronize()
            assert (sw.sum(0) - ref).abs().max().item() < tol, (splits, sw.shape)
            assert (pl.sum(0) - ref).abs().max().item() < tol, (splits, pl.shape)
            assert (sw.sum(0) - pl.sum(0)).abs().max().item() < tol
    finally:
        lib.vv_set_option(b"gemm_swap", 0)        # the library default


@pytest.mark.parametrize("M,N,K", [(128, 1536, 4608), (4096, 32, 128), (333, 256, 1024)])
def test_residual_epilogues(M, N, K):
    a, w = _mk(M, N, K, seed=4)
    acc = _ref(a, w)
    # decoder: x + gamma * (acc + bias), bf16 in/out (modular_vibevoice_tokenizer.py:940-942)
    bias = torch.randn(N, device="cuda") * 0.1
    gamma = torch.rand(N, device="cuda") * 0.4 + 0.2
    x = torch.randn(M, N, device="cuda").bfloat16()
    out = ops.linear(a, w, mode=_lib.GM_RESID_GAMMA_BF16, bias=bias, gamma=gamma, resid=x)
    torch.cuda.synchronize()
    r = x.float() + gamma * (acc + bias)
    assert ((out.float() - r).abs() <= 2e-3 + 8e-3 * r.abs()).all()
    # head: x + gate * acc, fp32 residual stream (modular_vibevoice_diffusion_head.py:160)
    xf = torch.randn(M, N, device="cuda")
    gate = torch.randn(M, N, device="cuda").bfloat16()
    out = ops.linear(a, w, mode=_lib.GM_RESID_GATE_F32, resid=xf, gate=gate)
    torch.cuda.synchronize()
    r = xf + gate.float() * acc
    assert (out - r).abs().max().item() < 5e-3


def test_strided_operands():
    # A and W taken as column slices of wider buffers (lda/ldw != K)
    g = torch.Generator().manual_seed(5)
    abuf = (torch.randn(200, 512, generator=g) * 0.5).bfloat16().cuda()
    wbuf = (torch.randn(384, 512, generator=g) * 0.05).bfloat16().cuda()
    a, w = abuf[:, 128:384], wbuf[:, 256:512]
    out = ops.linear(a, w, mode=_lib.GM_STORE_F32)
    torch.cuda.synchronize()
    assert (out - _ref(a, w)).abs().max().item() < 2e-3


def test_errors_are_loud():
    a, w = _mk(16, 64, 64)
    with pytest.raises(_lib.VVError):
        ops.linear(a[:, :60].contiguous(), w[:, :60].contiguous(), mode=_lib.GM_STORE_F32)  # K=60: ld not a multiple of 8
