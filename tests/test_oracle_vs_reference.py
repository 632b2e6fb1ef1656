"""Pins oracle/vv_oracle.py against the LIVE reference modules (imported from /root/reference
through oracle/ref_shim.py).  Runs only where the reference tree exists (this container);
on the GPU box the committed fixtures in tests/golden/ carry the same pins.
"""
import numpy as np
import pytest
import torch

from oracle import ref_shim
from oracle import vv_oracle as O
from vibevoice_b200 import synth

pytestmark = pytest.mark.skipif(not ref_shim.available(), reason="reference tree not present")

HEAD_SMALL = synth.HeadCfg(hidden=128, layers=2)


def _ref_head(cfg, W):
    ref = ref_shim.load()
    hc = ref.HeadConfig(hidden_size=cfg.hidden, latent_size=cfg.latent, head_layers=cfg.layers,
                        head_ffn_ratio=cfg.ffn_ratio, rms_norm_eps=cfg.eps)
    head = ref.Head(hc).eval()
    missing, unexpected = head.load_state_dict(W, strict=True)
    return head


def _ref_acoustic(cfg, W):
    ref = ref_shim.load()
    depths = "-".join(str(d) for d in reversed(cfg.depths))
    ac = ref.AcousticConfig(vae_dim=cfg.vae_dim, encoder_n_filters=cfg.n_filters,
                            decoder_n_filters=cfg.n_filters, encoder_ratios=list(cfg.ratios),
                            decoder_ratios=list(cfg.ratios), encoder_depths=depths,
                            layernorm_eps=cfg.eps)
    model = ref.Acoustic(ac).eval()
    res = model.load_state_dict({"decoder." + k: v for k, v in W.items()}, strict=False)
    assert not res.unexpected_keys
    assert all(k.startswith("encoder.") for k in res.missing_keys)
    return model


@pytest.mark.parametrize("S", [1, 2, 5, 10, 20])
def test_schedule_tables_match_reference(S):
    sch = ref_shim.make_scheduler()
    sch.set_timesteps(S)
    ts, sg = O.set_timesteps(S)
    assert np.array_equal(sch.timesteps.numpy(), ts)
    assert np.array_equal(sch.sigmas.numpy(), sg)


def test_schedule_known_answers():
    # SURVEY.md §8(c) pin 1 (probed on the live reference)
    ts, sg = O.set_timesteps(10)
    assert ts.tolist() == [999, 899, 799, 699, 599, 500, 400, 300, 200, 100]
    np.testing.assert_allclose(sg[1:4], [6.36459, 3.10373, 1.9807], rtol=1e-5)
    assert O.round_timesteps(ts, torch.bfloat16).tolist() == [1000, 900, 800, 700, 600, 500, 400, 300, 200, 100]
    ac = O.cosine_alphas_cumprod()
    assert abs(float(ac[0]) - 0.99995869) < 1e-7
    c = O.solver_coefficients(sg)
    np.testing.assert_allclose([c[0]["cx"], -c[0]["c0"]], [0.987881, -0.155166], rtol=1e-5)
    np.testing.assert_allclose([c[8]["cx"], -c[8]["c0"]], [0.52728, -0.486085], rtol=1e-5)
    assert c[9]["cx"] == 0.0 and c[9]["c0"] == 1.0


def test_head_forward_matches_reference():
    cfg = HEAD_SMALL
    W = synth.init_head(cfg, seed=7)
    head = _ref_head(cfg, W)
    g = torch.Generator().manual_seed(0)
    M = 6
    z = torch.randn(M, cfg.latent, generator=g)
    cond = torch.randn(M, cfg.hidden, generator=g)
    t = torch.tensor([999.0, 500.0, 3.0, 100.0, 0.0, 731.0])
    with torch.no_grad():
        want = head(z, t, cond)
    got = O.head_forward(W, z, t, cond, layers=cfg.layers, eps=cfg.eps, freq_dim=cfg.freq_dim)
    assert want.abs().max() > 0.05  # observable init: output is not trivially zero
    torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("algo", ["dpmsolver++", "sde-dpmsolver++"])
@pytest.mark.parametrize("S", [1, 2, 5, 10, 20])
def test_sample_speech_tokens_matches_reference(algo, S):
    cfg = HEAD_SMALL
    W = synth.init_head(cfg, seed=7)
    head = _ref_head(cfg, W)
    sch = ref_shim.make_scheduler(algo)
    n = 3
    cond, neg, noise = synth.synth_inputs(n, cfg.hidden, cfg.latent, seed=S)
    # the reference draws the SDE noise from the global CPU generator (dpm_solver.py:994-997)
    torch.manual_seed(99)
    want = ref_shim.ref_sample_speech_tokens(head, sch, cond, neg, 1.3, S, noise=noise)
    torch.manual_seed(99)
    sde = torch.stack([torch.randn(2 * n, cfg.latent) for _ in range(S)]) if algo.startswith("sde") else None
    got = O.sample_speech_tokens(W, cond, neg, 1.3, noise, steps=S, algo=algo, sde_noise=sde,
                                 t_dtype=torch.float32, layers=cfg.layers, eps=cfg.eps)
    torch.testing.assert_close(got, want, rtol=2e-4, atol=2e-5)


def test_reference_draws_noise_as_documented():
    """The shim's noise=None path is the verbatim reference draw: torch.randn(2n, L) from the
    global CPU generator; feeding the same tensor explicitly gives the same latent."""
    cfg = HEAD_SMALL
    W = synth.init_head(cfg, seed=7)
    head = _ref_head(cfg, W)
    cond, neg, _ = synth.synth_inputs(2, cfg.hidden, cfg.latent)
    torch.manual_seed(5)
    a = ref_shim.ref_sample_speech_tokens(head, ref_shim.make_scheduler(), cond, neg, 1.3, 5)
    torch.manual_seed(5)
    noise = torch.randn(4, cfg.latent)
    b = ref_shim.ref_sample_speech_tokens(head, ref_shim.make_scheduler(), cond, neg, 1.3, 5, noise=noise)
    assert torch.equal(a, b)


def test_cfg_scale_one_is_cond_branch():
    cfg = HEAD_SMALL
    W = synth.init_head(cfg, seed=7)
    cond, neg, noise = synth.synth_inputs(2, cfg.hidden, cfg.latent)
    a = O.sample_speech_tokens(W, cond, neg, 1.0, noise, steps=4, layers=cfg.layers)
    b = O.sample_speech_tokens(W, cond, torch.zeros_like(neg), 1.0, noise, steps=4, layers=cfg.layers)
    torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6)


def test_decoder_non_streaming_matches_reference():
    cfg = synth.DEC_TINY
    W = synth.init_decoder(cfg, seed=3)
    model = _ref_acoustic(cfg, W)
    g = torch.Generator().manual_seed(1)
    for T in (1, 3, 7):
        lat = torch.randn(2, cfg.vae_dim, T, generator=g)
        with torch.no_grad():
            want = model.decode(lat)
        got = O.decode(W, lat, cfg.depths, cfg.ratios, cfg.vae_dim, cfg.eps)
        assert want.shape == (2, 1, cfg.hop * T)
        assert want.abs().max() > 0.1
        torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-5)
    # [n, T, vae] input is permuted (modular_vibevoice_tokenizer.py:1102-1105)
    lat = torch.randn(2, 5, cfg.vae_dim, generator=g)
    with torch.no_grad():
        want = model.decode(lat)
    torch.testing.assert_close(O.decode(W, lat, cfg.depths, cfg.ratios, cfg.vae_dim, cfg.eps), want,
                               rtol=1e-4, atol=1e-5)


def test_decoder_streaming_matches_reference_cache():
    """Streaming with the reference's own cache class: all rows start together (so the
    late-joiner quirk of cache.get, SURVEY.md §8(c) 2C, does not fire), then a ragged subset,
    then set_to_zero on one row mid-stream."""
    ref = ref_shim.load()
    cfg = synth.DEC_TINY
    W = synth.init_decoder(cfg, seed=3)
    model = _ref_acoustic(cfg, W)
    g = torch.Generator().manual_seed(2)
    n = 3
    cache = ref.Cache()
    state = O.DecoderState(n)
    plan = [([0, 1, 2], 1), ([0, 1, 2], 1), ([0, 2], 1), ([0, 1, 2], 2), ("zero", [1]), ([0, 1, 2], 1), ([1], 1)]
    for step in plan:
        if step[0] == "zero":
            idx = torch.tensor(step[1])
            cache.set_to_zero(idx)
            state.set_to_zero(idx)
            continue
        rows, T = step
        idx = torch.tensor(rows)
        lat = torch.randn(len(rows), T, cfg.vae_dim, generator=g)
        with torch.no_grad():
            want = model.decode(lat, cache=cache, sample_indices=idx, use_cache=True)
        got = O.decode(W, lat, cfg.depths, cfg.ratios, cfg.vae_dim, cfg.eps, state=state, idx=idx)
        torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-5)


QUIRK_PLAN = [([0, 1], 1), ([0, 1], 1), ([0, 1, 2], 1), ([0, 1, 2], 1), ([1, 3], 2), ("zero", [0]), ([0, 1, 2, 3], 1)]


def test_late_joiner_quirk_matches_reference_cache():
    """a row that joins a running stream has no cache entry, so the reference's
    cache.get returns None (:204-207) and SConv1d / SConvTranspose1d restart EVERY row of that call from zero history
    (:337-346, :488-492).  DecoderState(strict_reference_quirk=True) reproduces it against the reference's own cache
    class; the default (per-row) state differs on exactly those calls."""
    ref = ref_shim.load()
    cfg = synth.DEC_TINY
    W = synth.init_decoder(cfg, seed=3)
    model = _ref_acoustic(cfg, W)
    g = torch.Generator().manual_seed(11)
    cache = ref.Cache()
    strict, sane = O.DecoderState(4, strict_reference_quirk=True), O.DecoderState(4)
    differs = []
    for step in QUIRK_PLAN:
        if step[0] == "zero":
            idx = torch.tensor(step[1])
            cache.set_to_zero(idx)
            strict.set_to_zero(idx)
            sane.set_to_zero(idx)
            continue
        rows, T = step
        idx = torch.tensor(rows)
        lat = torch.randn(len(rows), T, cfg.vae_dim, generator=g)
        with torch.no_grad():
            want = model.decode(lat, cache=cache, sample_indices=idx, use_cache=True)
        got = O.decode(W, lat, cfg.depths, cfg.ratios, cfg.vae_dim, cfg.eps, state=strict, idx=idx)
        torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-5)
        other = O.decode(W, lat, cfg.depths, cfg.ratios, cfg.vae_dim, cfg.eps, state=sane, idx=idx)
        differs.append(not torch.allclose(other, want, rtol=1e-3, atol=1e-4))
    # calls 3 and 5 (0-based 2 and 4) bring in a new row: there, and afterwards (the histories have diverged), the
    # per-row semantics give other audio than the reference
    assert differs[:2] == [False, False] and differs[2] and differs[4]


def test_streaming_equals_non_streaming_reference_and_oracle():
    cfg = synth.DEC_TINY
    W = synth.init_decoder(cfg, seed=3)
    g = torch.Generator().manual_seed(4)
    lat = torch.randn(2, cfg.vae_dim, 6, generator=g)
    full = O.decode(W, lat, cfg.depths, cfg.ratios, cfg.vae_dim, cfg.eps)
    state = O.DecoderState(2)
    idx = torch.arange(2)
    chunks = [O.decode(W, lat[:, :, t:t + 1], cfg.depths, cfg.ratios, cfg.vae_dim, cfg.eps, state=state, idx=idx)
              for t in range(6)]
    torch.testing.assert_close(torch.cat(chunks, dim=-1), full, rtol=1e-4, atol=1e-5)


# ---- semantic tokenizer encoder + connectors ----

def test_encoder_non_streaming_matches_reference():
    cfg = synth.ENC_TINY
    W = synth.init_encoder(cfg, seed=3)
    model = ref_shim.make_semantic(cfg, W)
    g = torch.Generator().manual_seed(1)
    for T in (1, 3, 6):
        audio = torch.randn(2, 1, cfg.hop * T, generator=g)
        with torch.no_grad():
            want = model.encode(audio).mean
        got = O.encode(W, audio, cfg.depths, cfg.ratios, cfg.eps)
        assert want.shape == (2, T, cfg.dim)
        assert want.abs().max() > 0.1
        torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-5)


def test_encoder_streaming_matches_reference_cache():
    """The frame loop's call (modeling_vibevoice_inference.py:673-679): one hop of audio per step through the
    reference's own cache class, a ragged row subset, and set_to_zero mid-stream."""
    ref = ref_shim.load()
    cfg = synth.ENC_TINY
    W = synth.init_encoder(cfg, seed=3)
    model = ref_shim.make_semantic(cfg, W)
    g = torch.Generator().manual_seed(2)
    n = 3
    cache = ref.Cache()
    state = O.DecoderState(n)
    plan = [([0, 1, 2], 1), ([0, 1, 2], 1), ([0, 2], 1), ([0, 1, 2], 2), ("zero", [1]), ([0, 1, 2], 1), ([1], 1)]
    for step in plan:
        if step[0] == "zero":
            idx = torch.tensor(step[1])
            cache.set_to_zero(idx)
            state.set_to_zero(idx)
            continue
        rows, T = step
        idx = torch.tensor(rows)
        audio = torch.randn(len(rows), 1, cfg.hop * T, generator=g)
        with torch.no_grad():
            want = model.encode(audio, cache=cache, sample_indices=idx, use_cache=True).mean
        got = O.encode(W, audio, cfg.depths, cfg.ratios, cfg.eps, state=state, idx=idx)
        torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-5)


def test_encoder_full_size_one_frame_matches_reference():
    """The real architecture (32..2048 channels, hop 3200, 128 dims), one frame, n = 1."""
    cfg = synth.ENC_FULL
    W = synth.init_encoder(cfg, seed=5)
    model = ref_shim.make_semantic(cfg, W)
    audio = torch.randn(1, 1, cfg.hop, generator=torch.Generator().manual_seed(9))
    with torch.no_grad():
        want = model.encode(audio).mean
    got = O.encode(W, audio, cfg.depths, cfg.ratios, cfg.eps)
    assert want.shape == (1, 1, 128)
    torch.testing.assert_close(got, want, rtol=2e-4, atol=2e-5)


@pytest.mark.parametrize("dims", [(16, 64), (64, 96)])
def test_connector_matches_reference(dims):
    cfg = synth.ConnCfg(input_dim=dims[0], output_dim=dims[1])
    W = synth.init_connector(cfg, seed=2)
    m = ref_shim.make_connector(cfg, W)
    x = torch.randn(3, 1, cfg.input_dim, generator=torch.Generator().manual_seed(3))
    with torch.no_grad():
        want = m(x)
    torch.testing.assert_close(O.speech_connector(W, x, cfg.eps), want, rtol=1e-4, atol=1e-5)


# ---- voice-prompt prefill ----

@pytest.mark.parametrize("L", [1, 5, 11, 12, 13, 29, 50, 96])
def test_prefill_encoder_any_length_matches_reference(L):
    """acoustic_tokenizer.encode on a prompt whose length is not a whole number of frames
    (modeling_vibevoice_inference.py:154): every strided conv completes its last window with zeros."""
    cfg = synth.ACOUSTIC_ENC_TINY
    W = synth.init_encoder(cfg, seed=7)
    model = ref_shim.make_acoustic_encoder(cfg, W)
    audio = torch.randn(2, 1, L, generator=torch.Generator().manual_seed(L))
    with torch.no_grad():
        out = model.encode(audio)
    got = O.encode(W, audio, cfg.depths, cfg.ratios, cfg.eps)
    assert out.std == 0.5
    assert out.mean.shape == (2, -(-L // cfg.hop), cfg.dim)
    torch.testing.assert_close(got, out.mean, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("dist", ["fix", "gaussian", "none"])
def test_sample_latents_matches_reference_draw_for_draw(dist, dtype):
    """VibeVoiceTokenizerEncoderOutput.sample (:970-990): same generator state -> identical bits, which also pins
    the ORDER of the draws the host mirror must reproduce ('gaussian': randn(batch) first, then randn_like(mean))."""
    mean = torch.randn(3, 5, 16, generator=torch.Generator().manual_seed(4)).to(dtype)
    torch.manual_seed(11)
    want, want_std = ref_shim.encoder_output(mean, 0.5).sample(dist_type=dist)
    torch.manual_seed(11)
    row_noise = torch.randn(3, dtype=dtype) if dist == "gaussian" else None
    noise = torch.randn_like(mean) if dist in ("fix", "gaussian") else None
    got, got_std = O.sample_latents(mean, 0.5, dist, noise, row_noise)
    assert got.dtype == dtype
    assert torch.equal(got, want)
    if dist == "gaussian":
        assert torch.equal(got_std, want_std)
    else:
        assert got_std == want_std == 0.5


def test_acoustic_features_formula():
    # (latents + bias) * scale with 0-dim buffers keeps bf16 (modeling_vibevoice_inference.py:159)
    lat = torch.randn(2, 3, 8).to(torch.bfloat16)
    b, s = torch.tensor(-0.0703), torch.tensor(0.1962)
    got = O.acoustic_features(lat, b, s)
    assert got.dtype == torch.bfloat16
    assert torch.equal(got, (lat + b) * s)
