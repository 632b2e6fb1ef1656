"""vibevoice_b200.dropin: the three reference callables are rebound with identical signatures and
the cache generate() constructs is adopted.  The GPU box has no reference tree, so the GPU test
drives install() with a structural stand-in of the model object (same attribute names, same
state-dict names, same config fields); the CPU test checks the signatures against the LIVE
reference classes and that a CPU model is refused (no fallback)."""
import inspect
import math
import types

import pytest
import torch
import torch.nn as nn

from oracle import ref_shim
from oracle import vv_oracle as O
from vibevoice_b200 import synth
from vibevoice_b200._lib import VVError

HC = synth.HeadCfg(hidden=128, layers=2)
DC = synth.DecCfg(vae_dim=HC.latent, n_filters=8, ratios=(2, 2, 2), depths=(2, 1, 1, 2))


def _tree(W, device):
    """nn.Module tree whose state_dict() has exactly the names/values of W."""
    root = nn.Module()
    for name, val in W.items():
        mod = root
        parts = name.split(".")
        for p in parts[:-1]:
            if not hasattr(mod, p):
                mod.add_module(p, nn.Module())
            mod = getattr(mod, p)
        mod.register_parameter(parts[-1], nn.Parameter(val.clone().to(device), requires_grad=False))
    return root


class _RefCache:
    """Same surface as VibeVoiceTokenizerStreamingCache (modular_vibevoice_tokenizer.py:193-256)."""

    def __init__(self):
        self.cache = {}
        self.zeroed = []

    def set_to_zero(self, sample_indices):
        self.zeroed.append(sample_indices.tolist())

    def clear(self, layer_id=None, sample_indices=None):
        self.cache.clear()


def _fake_model(device, steps=4, algo="dpmsolver++"):
    Wh, Wd = synth.init_head(HC, seed=5), synth.init_decoder(DC, seed=5)
    head = _tree(Wh, device)
    head.config = types.SimpleNamespace(hidden_size=HC.hidden, head_ffn_ratio=HC.ffn_ratio, latent_size=HC.latent,
                                        head_layers=HC.layers, rms_norm_eps=HC.eps)
    ac = nn.Module()
    ac.add_module("decoder", _tree(Wd, device))
    ac.decoder.n_filters, ac.decoder.ratios, ac.decoder.depths = DC.n_filters, list(DC.ratios), list(DC.depths)
    ac.config = types.SimpleNamespace(vae_dim=DC.vae_dim, layernorm_eps=DC.eps)
    core = types.SimpleNamespace(prediction_head=head, acoustic_tokenizer=ac,
                                 noise_scheduler=types.SimpleNamespace(config=types.SimpleNamespace(algorithm_type=algo)))
    model = types.SimpleNamespace(model=core, config=types.SimpleNamespace(acoustic_vae_dim=HC.latent),
                                  ddpm_inference_steps=steps)
    return model, Wh, Wd


EC = synth.EncCfg(dim=16, n_filters=8, ratios=(2, 2, 2), depths=(2, 1, 1, 2))     # hop 8 == DC.hop
CA = synth.ConnCfg(input_dim=HC.latent, output_dim=HC.hidden)
CS = synth.ConnCfg(input_dim=EC.dim, output_dim=HC.hidden)


def _add_frame_tail(model, device):
    """semantic_tokenizer / acoustic_connector / semantic_connector with the reference's attribute names."""
    We, Wa, Ws = synth.init_encoder(EC, seed=5), synth.init_connector(CA, seed=5, tag="a"), synth.init_connector(CS, seed=5, tag="s")
    st = nn.Module()
    st.add_module("encoder", _tree(We, device))
    st.encoder.n_filters, st.encoder.ratios, st.encoder.depths = EC.n_filters, list(EC.ratios), list(EC.depths)
    st.config = types.SimpleNamespace(vae_dim=EC.dim, layernorm_eps=EC.eps)
    model.model.semantic_tokenizer = st
    for name, W, cfg in (("acoustic_connector", Wa, CA), ("semantic_connector", Ws, CS)):
        m = _tree(W, device)
        m.fc1.in_features, m.fc1.out_features = cfg.input_dim, cfg.output_dim
        m.norm.variance_epsilon = cfg.eps
        setattr(model.model, name, m)
    return We, Wa, Ws


def _snr(got, ref):
    return 10 * math.log10(ref.double().pow(2).sum().item() / max((got.double() - ref.double()).pow(2).sum().item(), 1e-30))


@pytest.mark.gpu
def test_install_on_structural_stand_in():
    from vibevoice_b200 import dropin

    model, Wh, Wd = _fake_model("cuda")
    inst = dropin.install(model, max_batch=4)
    n = 3
    cond, neg, _ = synth.synth_inputs(n, HC.hidden, HC.latent, seed=2)
    torch.manual_seed(123)
    lat = model.sample_speech_tokens(cond.cuda().bfloat16(), neg.cuda().bfloat16(), cfg_scale=1.3)
    assert lat.shape == (n, HC.latent) and lat.dtype == torch.bfloat16
    torch.manual_seed(123)
    noise = torch.randn(2 * n, HC.latent)     # the draw install() makes, as the reference does (:716)
    want = O.sample_speech_tokens(Wh, cond.bfloat16().float(), neg.bfloat16().float(), 1.3, noise, steps=4,
                                  t_dtype=torch.bfloat16, layers=HC.layers)
    rel = ((lat.float().cpu() - want).norm() / want.norm()).item()
    assert rel < 5e-2, rel
    # prediction_head.forward
    z = torch.randn(4, HC.latent)
    t = torch.tensor([999.0, 500.0, 3.0, 100.0])
    c4 = torch.randn(4, HC.hidden)
    got = model.model.prediction_head.forward(z.cuda(), t.cuda(), c4.cuda()).float().cpu()
    ref = O.head_forward(Wh, z, t, c4, layers=HC.layers)
    assert ((got - ref).norm() / ref.norm()).item() < 3e-2
    # decode: streaming through a reference-style cache object + mirrored set_to_zero
    cache = _RefCache()
    state = O.DecoderState(4)
    idx = torch.arange(n, device="cuda")
    frames = torch.randn(n, 3, HC.latent)
    for f in range(3):
        if f == 2:
            cache.set_to_zero(torch.tensor([1]))
            state.set_to_zero(torch.tensor([1]))
        w = model.model.acoustic_tokenizer.decode(frames[:, f:f + 1].cuda(), cache=cache, sample_indices=idx,
                                                  use_cache=True, debug=False)
        r = O.decode(Wd, frames[:, f:f + 1], DC.depths, DC.ratios, DC.vae_dim, state=state, idx=idx.cpu())
        assert w.shape == (n, 1, DC.hop) and _snr(w.float().cpu(), r) > 30
    assert cache.zeroed == [[1]]              # the reference's own bookkeeping still ran
    off = model.model.acoustic_tokenizer.decode(frames.cuda())
    assert _snr(off.float().cpu(), O.decode(Wd, frames, DC.depths, DC.ratios, DC.vae_dim)) > 30
    inst.uninstall()
    assert not hasattr(model, "sample_speech_tokens")       # instance rebinding removed


@pytest.mark.gpu
def test_install_frame_tail_on_structural_stand_in():
    """One frame iteration past the decoder: audio chunk -> semantic encode (own cache) -> both connectors ->
    summed embedding, exactly the calls of modeling_vibevoice_inference.py:673-684."""
    from vibevoice_b200 import dropin

    model, Wh, Wd = _fake_model("cuda")
    We, Wa, Ws = _add_frame_tail(model, "cuda")
    inst = dropin.install(model, max_batch=4, frame_tail=True)
    n = 3
    idx = torch.arange(n, device="cuda")
    acoustic_cache, semantic_cache = _RefCache(), _RefCache()
    dstate, estate = O.DecoderState(4), O.DecoderState(4)
    frames = torch.randn(n, 2, HC.latent)
    for f in range(2):
        lat = frames[:, f:f + 1]
        chunk = model.model.acoustic_tokenizer.decode(lat.cuda(), cache=acoustic_cache, sample_indices=idx, use_cache=True)
        feat = model.model.semantic_tokenizer.encode(chunk, cache=semantic_cache, sample_indices=idx, use_cache=True,
                                                     debug=False).mean
        emb = model.model.acoustic_connector.forward(lat.cuda()) + model.model.semantic_connector.forward(feat)
        r_chunk = O.decode(Wd, lat, DC.depths, DC.ratios, DC.vae_dim, state=dstate, idx=idx.cpu())
        r_feat = O.encode(We, r_chunk, EC.depths, EC.ratios, EC.eps, state=estate, idx=idx.cpu())
        r_emb = O.speech_connector(Wa, lat, CA.eps) + O.speech_connector(Ws, r_feat, CS.eps)
        assert feat.shape == (n, 1, EC.dim) and emb.shape == (n, 1, HC.hidden)
        assert _snr(feat.float().cpu(), r_feat) > 28, _snr(feat.float().cpu(), r_feat)
        assert _snr(emb.float().cpu(), r_emb) > 28, _snr(emb.float().cpu(), r_emb)
    inst.uninstall()


AC = synth.ACOUSTIC_ENC_TINY


@pytest.mark.gpu
def test_install_prefill_on_structural_stand_in():
    """_process_speech_inputs (modeling_vibevoice_inference.py:150-177): prompt audio -> encode -> sample -> bias/scale
    -> acoustic connector -> masked rows, for both speech types, with the reference's draws."""
    from vibevoice_b200 import dropin

    model, Wh, Wd = _fake_model("cuda")
    _, Wa, _ = _add_frame_tail(model, "cuda")
    We = synth.init_encoder(synth.EncCfg(dim=HC.latent, n_filters=AC.n_filters, ratios=AC.ratios, depths=AC.depths), seed=9)
    ecfg = synth.EncCfg(dim=HC.latent, n_filters=AC.n_filters, ratios=AC.ratios, depths=AC.depths)
    at = model.model.acoustic_tokenizer
    at.add_module("encoder", _tree(We, "cuda"))
    at.encoder.n_filters, at.encoder.ratios, at.encoder.depths = ecfg.n_filters, list(ecfg.ratios), list(ecfg.depths)
    at.fix_std, at.std_dist_type = 0.5, "gaussian"
    at.encode = lambda *a, **k: (_ for _ in ()).throw(AssertionError("torch path used"))
    model.model.speech_bias_factor = torch.tensor(-0.0703125, device="cuda")
    model.model.speech_scaling_factor = torch.tensor(0.1962890625, device="cuda")
    inst = dropin.install(model, max_batch=4, prefill=True, prefill_rows=8)
    n, L = 2, 5 * ecfg.hop + 7
    T = -(-L // ecfg.hop)
    audio = torch.randn(n, L, generator=torch.Generator().manual_seed(1))
    masks = torch.tensor([[True] * T, [True] * (T - 2) + [False] * 2])
    torch.manual_seed(77)
    feats, connected = model._process_speech_inputs(audio.cuda().bfloat16(), masks.cuda())
    assert feats.shape == (n, T, HC.latent) and feats.dtype == torch.bfloat16
    assert connected.shape == (2 * T - 2, HC.hidden)
    torch.manual_seed(77)
    rn = torch.randn(n, device="cuda", dtype=torch.bfloat16).cpu()
    noise = torch.randn(n, T, HC.latent, device="cuda", dtype=torch.bfloat16).cpu()
    mean = O.encode(We, audio.bfloat16().float(), ecfg.depths, ecfg.ratios, ecfg.eps, emulate_bf16=True)
    x, _ = O.sample_latents(mean.bfloat16(), 0.5, "gaussian", noise, rn)
    want = O.acoustic_features(x, -0.0703125, 0.1962890625).float()
    assert _snr(feats.float().cpu(), want) > 28, _snr(feats.float().cpu(), want)
    wc = O.speech_connector(Wa, feats.float().cpu(), CA.eps)[masks]
    assert _snr(connected.float().cpu(), wc) > 28
    # speech_type "pt": latents given, only sample + bias/scale + connector
    torch.manual_seed(78)
    f2, c2 = model._process_speech_inputs(mean.bfloat16().cuda(), masks.cuda(), speech_type="pt")
    torch.manual_seed(78)
    rn = torch.randn(n, device="cuda", dtype=torch.bfloat16).cpu()
    noise = torch.randn(n, T, HC.latent, device="cuda", dtype=torch.bfloat16).cpu()
    x, _ = O.sample_latents(mean.bfloat16(), 0.5, "gaussian", noise, rn)
    assert torch.equal(f2.cpu(), O.acoustic_features(x, -0.0703125, 0.1962890625))
    with pytest.raises(NotImplementedError):
        model._process_speech_inputs(audio.cuda(), masks, speech_type="mel")
    inst.uninstall()
    assert not hasattr(model, "_process_speech_inputs")


def test_install_refuses_cpu_model():
    from vibevoice_b200 import dropin

    model, _, _ = _fake_model("cpu")
    with pytest.raises(VVError):
        dropin.install(model)


@pytest.mark.skipif(not ref_shim.available(), reason="reference tree not present")
def test_signatures_match_the_reference():
    """Parameter names/defaults of the rebinds == the reference methods they replace."""
    import sys

    ref = ref_shim.load()
    from vibevoice_b200 import dropin

    src = inspect.getsource(dropin.install)
    want_head = list(inspect.signature(ref.Head.forward).parameters)[1:]
    want_dec = inspect.signature(ref.Acoustic.decode).parameters
    assert want_head == ["noisy_images", "timesteps", "condition"]
    assert "def head_forward(noisy_images, timesteps, condition)" in src
    names = [k for k in want_dec if k != "self"]
    assert names == ["latents", "cache", "sample_indices", "use_cache", "debug"]
    assert "def decode(latents, cache=None, sample_indices=None, use_cache=False, debug=False)" in src
    assert [want_dec[k].default for k in names[1:]] == [None, None, False, False]
    # sample_speech_tokens lives in modeling_vibevoice_inference.py (needs the Qwen2 stack to import):
    # read its def line from the source file instead of importing the module.
    path = ref_shim.REFERENCE_ROOT + "/vibevoice/modular/modeling_vibevoice_inference.py"
    line = [l for l in open(path) if "def sample_speech_tokens" in l][0]
    assert "(self, condition, neg_condition, cfg_scale=3.0)" in line
    assert "def sample_speech_tokens(self, condition, neg_condition, cfg_scale=3.0)" in src


@pytest.mark.skipif(not ref_shim.available(), reason="reference tree not present")
def test_spec_extraction_on_live_reference_modules():
    """Everything from_module() does before the device copy, on the LIVE reference classes (this container only): the
    config attributes it reads exist with the meaning assumed, and the state dicts carry every weight name synth's
    ordered lists expect, with the right shapes.  Also: the scheduler VibeVoiceModel builds passes check_scheduler, and
    a scheduler configured otherwise is refused."""
    from vibevoice_b200 import dropin
    from vibevoice_b200.decoder import AcousticDecoderB200
    from vibevoice_b200.encoder import connector_spec_from_module, encoder_spec_from_module
    from vibevoice_b200.head import DiffusionHeadB200

    ref = ref_shim.load()
    # diffusion head
    Wh = synth.init_head(HC, seed=5)
    hc = ref.HeadConfig(hidden_size=HC.hidden, latent_size=HC.latent, head_layers=HC.layers, head_ffn_ratio=HC.ffn_ratio,
                        rms_norm_eps=HC.eps)
    head = ref.Head(hc).eval()
    head.load_state_dict(Wh, strict=True)
    W, cfg = DiffusionHeadB200.spec_from_module(head)
    assert cfg == HC
    got = synth.ordered_weights(synth.head_param_shapes(cfg), W)
    assert len(got) == len(Wh) and all(torch.equal(a, Wh[k]) for a, k in zip(got, synth.head_param_shapes(cfg)))
    # acoustic tokenizer: decoder half and encoder half (+ the sampling attributes the prefill path reads)
    Wd = synth.init_decoder(DC, seed=5)
    ac = ref.AcousticConfig(vae_dim=DC.vae_dim, encoder_n_filters=DC.n_filters, decoder_n_filters=DC.n_filters,
                            encoder_ratios=list(DC.ratios), decoder_ratios=list(DC.ratios),
                            encoder_depths="-".join(str(d) for d in reversed(DC.depths)), layernorm_eps=DC.eps)
    acoustic = ref.Acoustic(ac).eval()
    acoustic.load_state_dict({"decoder." + k: v for k, v in Wd.items()}, strict=False)
    W, cfg = AcousticDecoderB200.spec_from_module(acoustic)
    assert cfg == DC
    got = synth.ordered_weights(synth.dec_param_shapes(cfg), W, what="decoder weight")
    assert all(torch.equal(a, Wd[k]) for a, k in zip(got, synth.dec_param_shapes(cfg)))
    AEC = synth.ACOUSTIC_ENC_TINY
    Wae = synth.init_encoder(AEC, seed=5)
    aenc = ref_shim.make_acoustic_encoder(AEC, Wae, fix_std=0.5, std_dist_type="gaussian")
    W, cfg = encoder_spec_from_module(aenc)
    assert cfg == AEC and aenc.fix_std == 0.5 and aenc.std_dist_type == "gaussian"
    synth.ordered_weights(synth.enc_param_shapes(cfg), W, what="encoder weight")
    # semantic tokenizer
    We = synth.init_encoder(EC, seed=5)
    W, cfg = encoder_spec_from_module(ref_shim.make_semantic(EC, We))
    assert cfg == EC
    got = synth.ordered_weights(synth.enc_param_shapes(cfg), W, what="encoder weight")
    assert all(torch.equal(a, We[k]) for a, k in zip(got, synth.enc_param_shapes(cfg)))
    # connectors
    for c, tag in ((CA, "a"), (CS, "s")):
        Wc = synth.init_connector(c, seed=5, tag=tag)
        W, cfg = connector_spec_from_module(ref_shim.make_connector(c, Wc))
        assert cfg == c
        synth.ordered_weights(synth.conn_param_shapes(cfg), W, what="connector weight")
    # a weight that is missing or mis-shaped is refused before any device work
    bad = dict(Wh)
    bad.pop("final_layer.linear.weight")
    with pytest.raises(KeyError):
        synth.ordered_weights(synth.head_param_shapes(HC), bad)
    bad = dict(Wh, **{"cond_proj.weight": torch.zeros(3, 3)})
    with pytest.raises(ValueError):
        synth.ordered_weights(synth.head_param_shapes(HC), bad)
    # scheduler configuration: the reference's own passes, anything this library does not implement is refused
    dropin.check_scheduler(ref_shim.make_scheduler("dpmsolver++"))
    dropin.check_scheduler(ref_shim.make_scheduler("sde-dpmsolver++"))
    for kw in ({"prediction_type": "epsilon"}, {"beta_schedule": "linear"}, {"num_train_timesteps": 500}, {"solver_order": 3},
               {"algorithm_type": "dpmsolver"}):
        args = dict(num_train_timesteps=1000, beta_schedule="cosine", prediction_type="v_prediction")
        args.update(kw)
        try:
            sch = ref.Scheduler(**args)
        except Exception:
            sch = types.SimpleNamespace(config=types.SimpleNamespace(**args))
        with pytest.raises(VVError, match="not implemented"):
            dropin.check_scheduler(sch)


@pytest.mark.gpu
def test_generation_with_changing_row_count_never_synchronises():
    """generate() hands the acoustic path only the rows in diffusion mode, a count that changes from frame to frame
    (modeling_vibevoice_inference.py:588-590).  Under install() the head runs at one fixed shape and the decoder /
    encoder keep one program per row count: 40 frames with a different active-row count each make the library
    synchronise the host only while it meets a count for the first time (program build); further passes over the same
    counts: not once (vv_host_sync_count).  The padded head gives each active row what an unpadded call gives it (to bf16 round-off)."""
    from vibevoice_b200 import _lib, dropin
    from vibevoice_b200.head import DiffusionHeadB200

    lib = _lib.load()
    model, Wh, Wd = _fake_model("cuda", steps=4)
    inst = dropin.install(model, max_batch=8)
    assert inst.head.fixed_rows == 8
    cache = _RefCache()
    g = torch.Generator().manual_seed(7)
    counts = [(3 * f) % 8 + 1 for f in range(40)]          # 1..8, changing every frame
    plain = DiffusionHeadB200(Wh, HC, max_rows=16, model_dtype=torch.bfloat16)

    def run_pass(check_values):
        for f, k in enumerate(counts):
            rows = torch.randperm(8, generator=g)[:k].sort().values
            cond = torch.randn(k, HC.hidden, generator=g).cuda().bfloat16()
            neg = torch.randn(k, HC.hidden, generator=g).cuda().bfloat16()
            torch.manual_seed(1000 + f)
            lat = model.sample_speech_tokens(cond, neg, cfg_scale=1.3)
            assert lat.shape == (k, HC.latent)
            if check_values and f < 8:
                torch.manual_seed(1000 + f)
                want = plain.sample_speech_tokens(cond, neg, 1.3, num_steps=4, noise=torch.randn(2 * k, HC.latent))
                # same rows, same noise; the padded launch has another row count, so the GEMM planner may split K
                # differently: equal up to fp32 summation order, i.e. single bf16 ulps
                rel = ((lat.float() - want.float()).norm() / want.float().norm()).item()
                assert rel < 1e-2, (f, rel)
            wave = model.model.acoustic_tokenizer.decode(lat.unsqueeze(1), cache=cache, sample_indices=rows.cuda(), use_cache=True)
            assert wave.shape == (k, 1, DC.hop)

    run_pass(check_values=True)                             # program builds (one per distinct count and path) happen here
    torch.cuda.synchronize()
    s0 = lib.vv_host_sync_count()
    for _ in range(2):
        run_pass(check_values=False)
        torch.cuda.synchronize()
    assert lib.vv_host_sync_count() == s0, "the drop-in path synchronised the host in steady state"
    inst.uninstall()
