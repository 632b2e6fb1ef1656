"""GPU parity tests of the semantic-tokenizer encoder step and the SpeechConnector,
through the C ABI, against the CPU oracle.

Tolerances (features are O(1) activations, same budget as the decoder waveform): SNR >= 30 dB vs the fp32
oracle, >= 40 dB vs the oracle rounding to bf16 at the CUDA pipeline's storage points; connector rel-L2 <= 2e-2.
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


def _enc(cfg, seed=11, max_rows=64):
    from vibevoice_b200.encoder import SemanticEncoderB200

    W = synth.init_encoder(cfg, seed=seed)
    return W, SemanticEncoderB200(W, cfg, max_rows=max_rows)


@pytest.mark.parametrize("n,T", [(1, 1), (3, 2), (2, 5)])
def test_encoder_offline_vs_oracle_tiny(n, T):
    cfg = synth.ENC_TINY
    W, enc = _enc(cfg)
    audio = torch.randn(n, 1, cfg.hop * T, generator=torch.Generator().manual_seed(n * 10 + T))
    got = enc.encode(audio.cuda()).mean.float().cpu()
    assert got.shape == (n, T, cfg.dim)
    assert _snr(got, O.encode(W, audio, cfg.depths, cfg.ratios, cfg.eps)) >= 30.0
    assert _snr(got, O.encode(W, audio, cfg.depths, cfg.ratios, cfg.eps, emulate_bf16=True)) >= 40.0


def test_encoder_bf16_io_and_2d_audio():
    cfg = synth.ENC_TINY
    W, enc = _enc(cfg)
    audio = torch.randn(2, cfg.hop * 3, generator=torch.Generator().manual_seed(5))
    got = enc.encode(audio.cuda().bfloat16()).mean
    assert got.dtype == torch.bfloat16 and got.shape == (2, 3, cfg.dim)
    want = O.encode(W, audio.bfloat16().float(), cfg.depths, cfg.ratios, cfg.eps, emulate_bf16=True)
    assert _snr(got.float().cpu(), want) >= 35.0


def test_encoder_streaming_equals_offline_and_rows_independent():
    cfg = synth.ENC_TINY
    W, enc = _enc(cfg)
    n, F = 4, 6
    audio = torch.randn(n, 1, cfg.hop * F, generator=torch.Generator().manual_seed(3)).cuda()
    off = enc.encode(audio).mean.float()
    cache = enc.new_cache(n)
    idx = torch.arange(n, device="cuda")
    chunks = [enc.encode(audio[:, :, t * cfg.hop:(t + 1) * cfg.hop], cache=cache, sample_indices=idx, use_cache=True).mean
              for t in range(F)]
    st = torch.cat(chunks, dim=1).float()
    assert _snr(st.cpu(), off.cpu()) >= 45.0
    one = enc.encode(audio[2:3]).mean.float()          # row 2 alone == row 2 of the batch
    assert _snr(one.cpu(), off[2:3].cpu()) >= 45.0


def test_encoder_set_to_zero_and_ragged_subsets():
    cfg = synth.ENC_TINY
    W, enc = _enc(cfg)
    g = torch.Generator().manual_seed(8)
    n = 3
    cache = enc.new_cache(n)
    state = O.DecoderState(n)
    plan = [([0, 1, 2], 1), ([0, 2], 2), ("zero", [2]), ([2, 0], 1), ([1], 1), ([0, 1, 2], 1)]
    for step in plan:
        if step[0] == "zero":
            cache.set_to_zero(torch.tensor(step[1]))
            state.set_to_zero(torch.tensor(step[1]))
            continue
        rows, T = step
        audio = torch.randn(len(rows), 1, cfg.hop * T, generator=g)
        idx = torch.tensor(rows)
        got = enc.encode(audio.cuda(), cache=cache, sample_indices=idx.cuda(), use_cache=True).mean.float().cpu()
        want = O.encode(W, audio, cfg.depths, cfg.ratios, cfg.eps, state=state, idx=idx, emulate_bf16=True)
        assert _snr(got, want) >= 38.0, (step, _snr(got, want))


def test_encoder_full_size_frame_step():
    """The real architecture (32..2048 channels, hop 3200 -> 128 dims): two streamed frames, n = 2."""
    cfg = synth.ENC_FULL
    W, enc = _enc(cfg, seed=5, max_rows=8)
    n = 2
    audio = 0.5 * torch.randn(n, 1, cfg.hop * 2, generator=torch.Generator().manual_seed(9))
    cache = enc.new_cache(n)
    idx = torch.arange(n, device="cuda")
    got = torch.cat([enc.encode(audio[:, :, t * cfg.hop:(t + 1) * cfg.hop].cuda(), cache=cache, sample_indices=idx,
                                use_cache=True).mean for t in range(2)], dim=1).float().cpu()
    want = O.encode(W, audio, cfg.depths, cfg.ratios, cfg.eps)
    assert got.shape == (n, 2, 128)
    assert _snr(got, want) >= 30.0, _snr(got, want)
    off = enc.encode(audio.cuda()).mean.float().cpu()
    assert _snr(off, got) >= 40.0


def test_encoder_errors():
    cfg = synth.ENC_TINY
    W, enc = _enc(cfg, max_rows=4)
    with pytest.raises(ValueError):                          # streaming wants whole frames; offline takes any length
        enc.encode(torch.zeros(1, 1, cfg.hop + 1).cuda(), cache=enc.new_cache(2), sample_indices=torch.tensor([0]),
                   use_cache=True)
    assert enc.encode(torch.zeros(1, 1, cfg.hop + 1).cuda()).mean.shape == (1, 2, cfg.dim)
    with pytest.raises(ValueError):
        enc.encode(torch.zeros(1, 1, 0).cuda())
    with pytest.raises(Exception):
        enc.encode(torch.zeros(5, 1, cfg.hop).cuda())          # exceeds max_rows
    with pytest.raises(AssertionError):
        enc.encode(torch.zeros(1, 1, cfg.hop).cuda(), cache=enc.new_cache(2), use_cache=True)   # no sample_indices


@pytest.mark.parametrize("dims,M", [((64, 1536), 64), ((128, 1536), 7), ((16, 64), 3)])
def test_connector_vs_oracle(dims, M):
    from vibevoice_b200.encoder import SpeechConnectorB200

    cfg = synth.ConnCfg(input_dim=dims[0], output_dim=dims[1])
    W = synth.init_connector(cfg, seed=4)
    conn = SpeechConnectorB200(W, cfg, max_rows=64)
    x = torch.randn(M, 1, cfg.input_dim, generator=torch.Generator().manual_seed(M))
    got = conn(x.cuda()).float().cpu()
    want = O.speech_connector(W, x, cfg.eps)
    assert got.shape == want.shape
    rel = ((got - want).norm() / want.norm()).item()
    assert rel <= 2e-2, rel
    # the fused sum: acoustic_embed + semantic_embed (modeling_vibevoice_inference.py:684)
    add = torch.randn_like(want)
    got2 = conn(x.cuda(), addend=add.cuda()).float().cpu()
    rel2 = ((got2 - (want + add.bfloat16().float())).norm() / (want + add).norm()).item()
    assert rel2 <= 2e-2, rel2


def test_full_size_batch64_encoder_frame_steps_properties():
    """The frame loop's full size (batch 64, full semantic encoder, one frame per step; the last stage runs as one
    persistent launch): rows independent (same audio -> same features bit for bit in different rows and slots), reset
    == fresh, one row alone agrees with the batch, two spot rows meet the oracle gate."""
    cfg = synth.ENC_FULL
    W, enc = _enc(cfg, seed=7, max_rows=64)
    n, F = 64, 2
    g = torch.Generator().manual_seed(65)
    audio = 0.3 * torch.randn(n, F, cfg.hop, generator=g)
    audio[41] = audio[3]
    cache = enc.new_cache(n)
    slots = torch.randperm(n, generator=g)
    feats = [enc.encode(audio[:, f:f + 1].cuda(), cache=cache, sample_indices=slots.cuda(), use_cache=True).mean.float().cpu()
             for f in range(F)]
    out = torch.cat(feats, 1)
    assert out.shape == (n, F, cfg.dim) and bool(torch.isfinite(out).all())
    assert torch.equal(out[41], out[3])
    solo_cache = enc.new_cache(1)
    solo = torch.cat([enc.encode(audio[8:9, f:f + 1].cuda(), cache=solo_cache, sample_indices=torch.zeros(1, dtype=torch.long).cuda(),
                                 use_cache=True).mean.float().cpu() for f in range(F)], 1)
    assert _snr(solo, out[8:9]) >= 30.0, _snr(solo, out[8:9])
    cache.set_to_zero(slots[3:4])
    again = enc.encode(audio[:, 0:1].cuda(), cache=cache, sample_indices=slots.cuda(), use_cache=True).mean.float().cpu()
    assert torch.equal(again[3], feats[0][3])
    assert not torch.equal(again[4], feats[0][4])
    rows = torch.tensor([8, 50])
    state = O.DecoderState(2)
    want = torch.cat([O.encode(W, audio[rows, f:f + 1], cfg.depths, cfg.ratios, cfg.eps, state=state, idx=torch.arange(2))
                      for f in range(F)], 1)
    assert _snr(out[rows], want) >= 28.0, _snr(out[rows], want)
