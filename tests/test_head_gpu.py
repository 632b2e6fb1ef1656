"""Diffusion head + DPM-Solver++ loop on the GPU (through the C ABI) vs the CPU oracle.

Two comparisons per case:
  * vs oracle(emulate_bf16=True): same bf16 rounding points => tight tolerance; catches logic bugs.
  * vs oracle fp32 "truth": the north_star tolerance — rel-L2 of the final latent.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import vv_oracle as O  # noqa: E402  (checker only)
from vibevoice_b200 import synth  # noqa: E402
from vibevoice_b200.head import DiffusionHeadB200  # noqa: E402

SMALL = synth.HeadCfg(hidden=256, layers=2)


def _rel(a, b):
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


@pytest.fixture(scope="module")
def small_head():
    W = synth.init_head(SMALL, seed=11)
    return W, DiffusionHeadB200(W, SMALL, max_rows=1024)


@pytest.fixture(scope="module")
def full_head():
    W = synth.init_head(synth.HEAD_1P5B, seed=1234)
    return W, DiffusionHeadB200(W, synth.HEAD_1P5B, max_rows=1024)


@pytest.mark.parametrize("M", [1, 2, 6, 128, 130, 600])
def test_forward_small(small_head, M):
    W, head = small_head
    g = torch.Generator().manual_seed(M)
    z = torch.randn(M, SMALL.latent, generator=g)
    cond = torch.randn(M, SMALL.hidden, generator=g)
    t = torch.randint(0, 1000, (M,), generator=g).float()
    got = head.forward(z.cuda(), t.cuda(), cond.cuda()).float().cpu()
    emu = O.head_forward(W, z, t, cond, layers=SMALL.layers, eps=SMALL.eps, emulate_bf16=True)
    tru = O.head_forward(W, z, t, cond, layers=SMALL.layers, eps=SMALL.eps)
    assert tru.abs().max() > 0.05
    assert _rel(got, emu) < 1e-2, _rel(got, emu)
    assert _rel(got, tru) < 3e-2, _rel(got, tru)


def test_forward_bf16_io(small_head):
    W, head = small_head
    g = torch.Generator().manual_seed(3)
    M = 16
    z = torch.randn(M, SMALL.latent, generator=g).bfloat16()
    cond = torch.randn(M, SMALL.hidden, generator=g).bfloat16()
    t = torch.full((M,), 500.0)
    got = head.forward(z.cuda(), t.cuda().bfloat16(), cond.cuda())
    assert got.dtype == torch.bfloat16
    tru = O.head_forward(W, z.float(), t, cond.float(), layers=SMALL.layers, eps=SMALL.eps)
    assert _rel(got.float().cpu(), tru) < 3e-2


@pytest.mark.parametrize("algo", ["dpmsolver++", "sde-dpmsolver++"])
@pytest.mark.parametrize("S", [1, 2, 5, 10, 20])
def test_sample_small(small_head, algo, S):
    W, head = small_head
    n = 5
    cond, neg, noise = synth.synth_inputs(n, SMALL.hidden, SMALL.latent, seed=S)
    g = torch.Generator().manual_seed(S + 100)
    sde = torch.randn(S, 2 * n, SMALL.latent, generator=g) if algo.startswith("sde") else None
    got = head.sample_speech_tokens(cond.cuda(), neg.cuda(), 1.3, num_steps=S, algorithm_type=algo,
                                    noise=noise, sde_noise=None if sde is None else sde[:, :n].contiguous())
    got = got.float().cpu()
    kw = dict(steps=S, algo=algo, sde_noise=sde, t_dtype=torch.bfloat16, layers=SMALL.layers, eps=SMALL.eps)
    emu = O.sample_speech_tokens(W, cond, neg, 1.3, noise, emulate_bf16=True, **kw)
    tru = O.sample_speech_tokens(W, cond, neg, 1.3, noise, **kw)
    assert _rel(got, emu) < 2e-2, _rel(got, emu)
    assert _rel(got, tru) < 5e-2, _rel(got, tru)


def test_sample_full_1p5b_config1_and_batch64(full_head):
    """1.5B head, 10 steps, CFG 1.3 (BASELINE.json configs[0]/[1]); n = 1 and n = 64."""
    W, head = full_head
    cfg = synth.HEAD_1P5B
    for n in (1, 64):
        cond, neg, noise = synth.synth_inputs(n, cfg.hidden, cfg.latent)
        got = head.sample_speech_tokens(cond.cuda(), neg.cuda(), 1.3, num_steps=10, noise=noise).float().cpu()
        k = min(n, 4)  # the oracle costs seconds per row-step on CPU: check a slice of rows
        sl = slice(0, k)
        emu = O.sample_speech_tokens(W, cond[sl], neg[sl], 1.3, torch.cat([noise[:n][sl], noise[n:][sl]]),
                                     steps=10, emulate_bf16=True)
        tru = O.sample_speech_tokens(W, cond[sl], neg[sl], 1.3, torch.cat([noise[:n][sl], noise[n:][sl]]), steps=10)
        assert _rel(got[sl], emu) < 2e-2, (n, _rel(got[sl], emu))
        assert _rel(got[sl], tru) < 5e-2, (n, _rel(got[sl], tru))


def test_rows_are_independent(small_head):
    """Batch rows never mix (data-parallel sharding relies on it): row i of a batch of 7 equals a
    batch-of-1 run on row i."""
    W, head = small_head
    n = 7
    cond, neg, noise = synth.synth_inputs(n, SMALL.hidden, SMALL.latent, seed=5)
    full = head.sample_speech_tokens(cond.cuda(), neg.cuda(), 1.3, num_steps=5, algorithm_type="dpmsolver++",
                                     noise=noise).cpu()
    for i in (0, 3, 6):
        one = head.sample_speech_tokens(cond[i:i + 1].cuda(), neg[i:i + 1].cuda(), 1.3, num_steps=5,
                                        noise=torch.cat([noise[i:i + 1], noise[n + i:n + i + 1]])).cpu()
        assert _rel(one[0].float(), full[i].float()) < 2e-3


def test_cfg_one_ignores_negative(small_head):
    W, head = small_head
    cond, neg, noise = synth.synth_inputs(3, SMALL.hidden, SMALL.latent, seed=8)
    a = head.sample_speech_tokens(cond.cuda(), neg.cuda(), 1.0, num_steps=4, algorithm_type="dpmsolver++", noise=noise)
    b = head.sample_speech_tokens(cond.cuda(), torch.zeros_like(neg).cuda(), 1.0, num_steps=4, noise=noise)
    assert _rel(a.float(), b.float()) < 1e-5


def test_errors(small_head):
    from vibevoice_b200._lib import VVError
    W, head = small_head
    cond, neg, noise = synth.synth_inputs(600, SMALL.hidden, SMALL.latent)
    with pytest.raises(VVError):
        head.sample_speech_tokens(cond.cuda(), neg.cuda(), 1.3, num_steps=2, noise=noise)  # 2n > max_rows


@pytest.mark.parametrize("algo", ["dpmsolver++", "sde-dpmsolver++"])
def test_persistent_chain_matches_per_launch_path(full_head, algo):
    """vv_head_sample as ONE persistent cooperative launch (vv_chain.cu) against the per-op launch
    sequence it replaces: same kernels' arithmetic, only the split-K factors (fp32 summation order)
    differ."""
    from vibevoice_b200 import _lib

    W, head = full_head
    lib = _lib.load()
    cfg = synth.HEAD_1P5B
    outs = {}
    try:
        for n in (64, 3, 100, 200):  # 100 / 200 utterances = 200 / 400 rows: two / four 128-row tiles per GEMM phase, several rounds of row phases
            cond, neg, noise = synth.synth_inputs(n, cfg.hidden, cfg.latent, seed=n)
            g = torch.Generator().manual_seed(5)
            sde = torch.randn(10, n, cfg.latent, generator=g) if algo.startswith("sde") else None
            for chain in (1, 0):
                lib.vv_set_option(b"head_chain", chain)
                outs[chain] = head.sample_speech_tokens(cond.cuda(), neg.cuda(), 1.3, num_steps=10, algorithm_type=algo,
                                                        noise=noise, sde_noise=sde).float().cpu()
            assert torch.isfinite(outs[1]).all()
            assert _rel(outs[1], outs[0]) < 1e-2, (n, _rel(outs[1], outs[0]))
    finally:
        lib.vv_set_option(b"head_chain", 1)


def test_large_head_chain_and_per_launch_paths():
    """VibeVoice-Large head (hidden 3584, BASELINE.json configs[3]): the persistent kernel's staging
    layout and tile counts differ from the 1.5B head (56 column tiles, 2-way split-K, 2-tile weight
    chunks in the step-boundary phase).  Both paths against the bf16-emulating oracle and each other."""
    from vibevoice_b200 import _lib

    cfg = synth.HEAD_LARGE
    W = synth.init_head(cfg, seed=7)
    head = DiffusionHeadB200(W, cfg, max_rows=16)
    lib = _lib.load()
    n, S = 2, 3
    cond, neg, noise = synth.synth_inputs(n, cfg.hidden, cfg.latent, seed=3)
    outs = {}
    try:
        for chain in (1, 0):
            lib.vv_set_option(b"head_chain", 2 if chain else 0)     # 2: force the persistent kernel (auto prefers per-launch for H > 2048)
            outs[chain] = head.sample_speech_tokens(cond.cuda(), neg.cuda(), 1.3, num_steps=S, noise=noise).float().cpu()
    finally:
        lib.vv_set_option(b"head_chain", 1)
    assert torch.isfinite(outs[1]).all()
    assert _rel(outs[1], outs[0]) < 1e-2, _rel(outs[1], outs[0])
    emu = O.sample_speech_tokens(W, cond[:1], neg[:1], 1.3, torch.cat([noise[:n][:1], noise[n:][:1]]), steps=S,
                                 emulate_bf16=True)
    assert _rel(outs[1][:1], emu) < 2e-2, _rel(outs[1][:1], emu)
    assert _rel(outs[0][:1], emu) < 2e-2, _rel(outs[0][:1], emu)


def test_chain_program_cache_follows_batch_size_and_schedule(full_head):
    """A streaming server steps a varying subset of rows and may change the step count: every (n, S)
    gets its own persistent program; results must not depend on the order of the calls."""
    W, head = full_head
    cfg = synth.HEAD_1P5B
    ref = {}
    for n, S in [(5, 4), (2, 4), (5, 6)]:
        cond, neg, noise = synth.synth_inputs(n, cfg.hidden, cfg.latent, seed=10 * n + S)
        ref[(n, S)] = (cond, neg, noise,
                       head.sample_speech_tokens(cond.cuda(), neg.cuda(), 1.3, num_steps=S, noise=noise,
                                                 algorithm_type="dpmsolver++").float().cpu())
    for key in [(2, 4), (5, 6), (5, 4), (2, 4)]:
        cond, neg, noise, want = ref[key]
        got = head.sample_speech_tokens(cond.cuda(), neg.cuda(), 1.3, num_steps=key[1], noise=noise,
                                        algorithm_type="dpmsolver++").float().cpu()
        # same program, same inputs: every reduction in the solver loop runs in a fixed order (no atomics)
        assert torch.equal(got, want), (key, _rel(got, want))


def test_sample_is_bitwise_reproducible(full_head):
    """vv_head_sample twice on the same inputs (batch 64, 10 steps, both solvers, fp32 output): identical bits.
    Split-K partials are summed in a fixed order and the step-boundary phase's per-warp partial sums meet in warp
    order (vv_chain.cu, phase_tail), so nothing depends on timing."""
    W, head = full_head
    cfg = synth.HEAD_1P5B
    n = 64
    cond, neg, noise = synth.synth_inputs(n, cfg.hidden, cfg.latent, seed=77)
    for algo in ("dpmsolver++", "sde-dpmsolver++"):
        torch.manual_seed(5)
        sde = torch.randn(10, n, cfg.latent) if algo.startswith("sde") else None
        outs = []
        for _ in range(3):
            kw = {"sde_noise": sde} if sde is not None else {}
            outs.append(head.sample_speech_tokens(cond.cuda(), neg.cuda(), 1.3, num_steps=10, noise=noise.cuda(),
                                                  algorithm_type=algo, **kw).clone())
        torch.cuda.synchronize()
        assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2]), algo


@pytest.mark.parametrize("opt,val", [("head_tail_split", 1), ("chain_part_rowmajor", 0), ("head_fuse_tail", 0),
                                     ("chain_push_barrier", 1)])
def test_chain_program_variants_agree(full_head, opt, val):
    """The persistent head program's optional shapes (include/vvb200.h: the step boundary shared by two CTAs per row
    pair, plane-major split-K partials, the step boundary as five phases, unqualified proxy fences in the grid barrier)
    against the default program: same arithmetic, at most another fp32 summation order.  Program-shaping options apply
    to programs built afterwards, so the variant gets a head of its own."""
    from vibevoice_b200 import _lib

    W, head = full_head
    lib = _lib.load()
    cfg = synth.HEAD_1P5B
    prev = lib.vv_set_option(opt.encode(), val)
    try:
        other = DiffusionHeadB200(W, cfg, max_rows=128)
        for n, algo in ((64, "dpmsolver++"), (3, "sde-dpmsolver++")):
            cond, neg, noise = synth.synth_inputs(n, cfg.hidden, cfg.latent, seed=40 + n)
            g = torch.Generator().manual_seed(6)
            sde = torch.randn(10, n, cfg.latent, generator=g) if algo.startswith("sde") else None
            a = other.sample_speech_tokens(cond.cuda(), neg.cuda(), 1.3, num_steps=10, algorithm_type=algo, noise=noise,
                                           sde_noise=sde).float().cpu()
            a2 = other.sample_speech_tokens(cond.cuda(), neg.cuda(), 1.3, num_steps=10, algorithm_type=algo, noise=noise,
                                            sde_noise=sde).float().cpu()
            assert torch.isfinite(a).all() and torch.equal(a, a2)          # every variant is bitwise reproducible
            lib.vv_set_option(opt.encode(), prev)
            b = head.sample_speech_tokens(cond.cuda(), neg.cuda(), 1.3, num_steps=10, algorithm_type=algo, noise=noise,
                                          sde_noise=sde).float().cpu()
            lib.vv_set_option(opt.encode(), val)
            assert _rel(a, b) < 1e-2, (opt, n, _rel(a, b))
    finally:
        lib.vv_set_option(opt.encode(), prev)
