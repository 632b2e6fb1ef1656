"""GPU parity tests of the voice-prompt prefill through the C ABI, against the CPU oracle:
the acoustic tokenizer's encoder on prompts of ANY length (the last frame is completed with zeros stage by stage,
modular_vibevoice_tokenizer.py:127-133, :393-401), VibeVoiceTokenizerEncoderOutput.sample (:970-994) and the bias /
scale + connector of _process_speech_inputs (modeling_vibevoice_inference.py:150-164).

Tolerances: encoder SNR >= 30 dB vs the fp32 oracle and >= 40 dB vs the oracle rounding to bf16 at the CUDA pipeline's
storage points (same as tests/test_encoder_gpu.py); sampling and bias/scale are elementwise and must be BIT-EXACT.
"""
import math

import pytest
import torch

from oracle import vv_oracle as O
from vibevoice_b200 import synth

pytestmark = pytest.mark.gpu


def _snr(got, ref):
    err = (got.double() - ref.double()).pow(2).sum().item()
    return 10 * math.log10(ref.double().pow(2).sum().item() / max(err, 1e-30))


def _enc(cfg, seed=13, max_rows=64, **kw):
    from vibevoice_b200.encoder import AcousticEncoderB200

    W = synth.init_encoder(cfg, seed=seed)
    return W, AcousticEncoderB200(W, cfg, max_rows=max_rows, **kw)


@pytest.mark.parametrize("L", [1, 5, 11, 12, 13, 23, 24, 25, 50, 96, 487])
def test_prefill_any_length_vs_oracle_tiny(L):
    cfg = synth.ACOUSTIC_ENC_TINY                      # strides 2, 3, 2 (hop 12)
    W, enc = _enc(cfg)
    n = 3
    audio = torch.randn(n, 1, L, generator=torch.Generator().manual_seed(100 + L))
    out = enc.encode(audio.cuda())
    got = out.mean.float().cpu()
    assert out.std == 0.5
    assert got.shape == (n, -(-L // cfg.hop), cfg.dim)
    assert _snr(got, O.encode(W, audio, cfg.depths, cfg.ratios, cfg.eps)) >= 30.0
    assert _snr(got, O.encode(W, audio, cfg.depths, cfg.ratios, cfg.eps, emulate_bf16=True)) >= 40.0


def test_prefill_whole_frames_equals_offline_encode_bitwise():
    """n_samples a multiple of the hop: no masking anywhere, the same launches as vv_enc_encode."""
    from vibevoice_b200 import _lib
    from vibevoice_b200._lib import check, ptr

    cfg = synth.ACOUSTIC_ENC_TINY
    W, enc = _enc(cfg)
    audio = torch.randn(2, cfg.hop * 4, generator=torch.Generator().manual_seed(3)).cuda()
    a = enc.encode(audio).mean
    b = torch.empty_like(a)
    check(enc.lib.vv_enc_encode(enc._h, ptr(audio), 2, 4, ptr(b), _lib.VV_DTYPE_F32, None), "vv_enc_encode")
    torch.cuda.synchronize()
    assert torch.equal(a, b)


def test_prefill_time_chunking_and_tail_independence():
    """A handle with room for 2 frames per utterance per pass (chunked, history carried by the private cache)
    agrees with one that takes the prompt in one pass; samples past the prompt's end never leak in."""
    cfg = synth.ACOUSTIC_ENC_TINY
    L = cfg.hop * 9 + 5
    audio = torch.randn(2, 1, L, generator=torch.Generator().manual_seed(4))
    W, big = _enc(cfg, max_rows=64)
    _, small = _enc(cfg, max_rows=4)
    a = big.encode(audio.cuda()).mean.float().cpu()
    b = small.encode(audio.cuda()).mean.float().cpu()
    assert a.shape == b.shape == (2, 10, cfg.dim)
    assert _snr(b, a) >= 45.0
    assert _snr(b, O.encode(W, audio, cfg.depths, cfg.ratios, cfg.eps, emulate_bf16=True)) >= 40.0
    # the same prompt embedded in a longer buffer of garbage, sliced as a non-contiguous view
    buf = torch.full((2, 1, L + 40), 1e4)
    buf[:, :, :L] = audio
    c = big.encode(buf.cuda()[:, :, :L]).mean.float().cpu()
    assert torch.equal(c, a)
    # causal: a longer prompt with the same beginning has the same leading frames
    longer = torch.cat([audio, torch.randn(2, 1, 30, generator=torch.Generator().manual_seed(5))], dim=2)
    d = big.encode(longer.cuda()).mean.float().cpu()
    assert _snr(d[:, :9], a[:, :9]) >= 45.0


def test_prefill_bf16_io():
    cfg = synth.ACOUSTIC_ENC_TINY
    W, enc = _enc(cfg)
    audio = torch.randn(2, 1, 77, generator=torch.Generator().manual_seed(6))
    got = enc.encode(audio.cuda().bfloat16()).mean
    assert got.dtype == torch.bfloat16 and got.shape == (2, 7, cfg.dim)
    want = O.encode(W, audio.bfloat16().float(), cfg.depths, cfg.ratios, cfg.eps, emulate_bf16=True)
    assert _snr(got.float().cpu(), want) >= 35.0


def test_prefill_full_size_acoustic_encoder():
    """The real architecture (vae_dim 64, hop 3200, 32..2048 channels) on a prompt of 2.4 frames."""
    cfg = synth.ACOUSTIC_ENC_FULL
    W, enc = _enc(cfg, seed=5, max_rows=8)
    L = 2 * cfg.hop + 1234
    audio = 0.3 * torch.randn(1, 1, L, generator=torch.Generator().manual_seed(9))
    got = enc.encode(audio.cuda()).mean.float().cpu()
    assert got.shape == (1, 3, 64)
    want = O.encode(W, audio, cfg.depths, cfg.ratios, cfg.eps, emulate_bf16=True)
    assert _snr(got, want) >= 35.0, _snr(got, want)
    assert _snr(got, O.encode(W, audio, cfg.depths, cfg.ratios, cfg.eps)) >= 28.0


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("dist", ["fix", "gaussian", "none"])
def test_sample_draws_and_arithmetic_match_oracle_bitwise(dist, dtype):
    """EncoderOutputB200.sample makes the reference's draws in the reference's order on the mean's device, so a
    seeded run is reproducible against torch itself; the arithmetic is one kernel and bit-exact."""
    from vibevoice_b200.encoder import EncoderOutputB200

    mean = torch.randn(4, 7, 64, generator=torch.Generator().manual_seed(2)).to(dtype).cuda()
    torch.manual_seed(31)
    got, got_std = EncoderOutputB200(mean, 0.5).sample(dist)
    torch.manual_seed(31)
    rn = torch.randn(4, device="cuda", dtype=dtype) if dist == "gaussian" else None
    noise = torch.randn_like(mean) if dist in ("fix", "gaussian") else None
    want, want_std = O.sample_latents(mean.cpu(), 0.5, dist, noise.cpu() if noise is not None else None,
                                      rn.cpu() if rn is not None else None)
    assert got.dtype == dtype and torch.equal(got.cpu(), want)
    if dist == "gaussian":
        assert got_std.shape == (4, 1, 1) and torch.equal(got_std.cpu(), want_std)
    else:
        assert got_std == 0.5
    # fused bias/scale
    torch.manual_seed(31)
    feats = EncoderOutputB200(mean, 0.5).sample_features(dist, torch.tensor(-0.0703125), torch.tensor(0.1962890625))
    wantf = O.acoustic_features(want, -0.0703125, 0.1962890625)
    assert torch.equal(feats.cpu(), wantf)


def test_sampling_rejects_unknown_dist_like_the_reference():
    cfg = synth.ACOUSTIC_ENC_TINY
    _, enc = _enc(cfg)
    out = enc.encode(torch.randn(1, 1, 20).cuda())
    with pytest.raises(ValueError):
        enc.sampling(out, dist_type="laplace")
    x, std = enc.sampling(out)                       # default: the tokenizer's std_dist_type ('gaussian')
    assert x.shape == out.mean.shape and std.shape == (1, 1, 1)


def test_connector_row_chunking():
    from vibevoice_b200.encoder import SpeechConnectorB200

    cfg = synth.ConnCfg(input_dim=16, output_dim=64)
    W = synth.init_connector(cfg, seed=2)
    small, big = SpeechConnectorB200(W, cfg, max_rows=8), SpeechConnectorB200(W, cfg, max_rows=64)
    x = torch.randn(3, 9, 16, generator=torch.Generator().manual_seed(1)).cuda()
    a, b = small(x), big(x)
    assert a.shape == (3, 9, 64)
    rel = ((a.float() - b.float()).norm() / b.float().norm()).item()
    assert rel <= 1e-2, rel
    want = O.speech_connector(W, x.cpu(), cfg.eps)
    assert ((a.float().cpu() - want).norm() / want.norm()).item() <= 2e-2
