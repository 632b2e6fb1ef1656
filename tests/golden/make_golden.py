#!/usr/bin/env python
"""Generates tests/golden/*.npz from the LIVE reference (/root/reference), CPU fp32.

Run in the build container only (the reference tree does not exist on the GPU box):

    python tests/golden/make_golden.py

The reference ships no golden vectors for this path, so these fixtures are
outputs of the reference's own modules — VibeVoiceDiffusionHead, DPMSolverMultistepScheduler,
VibeVoiceAcousticTokenizerModel.decode, VibeVoiceTokenizerStreamingCache — on seeded inputs
with the deterministic "observable" init of vibevoice_b200/synth.py (weights are re-derived
from (seed, name) by the tests; only inputs-by-seed and reference OUTPUTS are stored).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_shim  # noqa: E402
from vibevoice_b200 import synth  # noqa: E402

HEAD_SMALL = synth.HeadCfg(hidden=128, layers=2)
HEAD_SEED = 7
DEC_SEED = 3
# streaming plan shared with tests/test_golden.py: (rows, T) or ("zero", rows)
STREAM_PLAN = [([0, 1, 2], 1), ([0, 1, 2], 1), ([0, 2], 1), ([0, 1, 2], 2), ("zero", [1]), ([0, 1, 2], 1), ([1], 1)]


def ref_head(cfg, W):
    ref = ref_shim.load()
    hc = ref.HeadConfig(hidden_size=cfg.hidden, latent_size=cfg.latent, head_layers=cfg.layers,
                        head_ffn_ratio=cfg.ffn_ratio, rms_norm_eps=cfg.eps)
    head = ref.Head(hc).eval()
    head.load_state_dict(W, strict=True)
    return head


def ref_acoustic(cfg, W):
    ref = ref_shim.load()
    depths = "-".join(str(d) for d in reversed(cfg.depths))
    ac = ref.AcousticConfig(vae_dim=cfg.vae_dim, encoder_n_filters=cfg.n_filters, decoder_n_filters=cfg.n_filters,
                            encoder_ratios=list(cfg.ratios), decoder_ratios=list(cfg.ratios), encoder_depths=depths,
                            layernorm_eps=cfg.eps)
    model = ref.Acoustic(ac).eval()
    model.load_state_dict({"decoder." + k: v for k, v in W.items()}, strict=False)
    return model


def main():
    assert ref_shim.available(), "reference tree not found"
    out = {}
    # ---- schedule tables (set_timesteps, schedule/dpm_solver.py:321-423) ----
    for S in (1, 2, 5, 10, 20):
        sch = ref_shim.make_scheduler()
        sch.set_timesteps(S)
        out[f"sched_ts_{S}"] = sch.timesteps.numpy().astype(np.int64)
        out[f"sched_sigmas_{S}"] = sch.sigmas.numpy().astype(np.float32)
    out["alphas_cumprod"] = ref_shim.make_scheduler().alphas_cumprod.numpy().astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "schedule.npz"), **out)

    # ---- head forward + sample_speech_tokens (small head, fp32) ----
    out = {}
    cfg = HEAD_SMALL
    W = synth.init_head(cfg, seed=HEAD_SEED)
    head = ref_head(cfg, W)
    g = torch.Generator().manual_seed(0)
    z = torch.randn(6, cfg.latent, generator=g)
    cond = torch.randn(6, cfg.hidden, generator=g)
    t = torch.tensor([999.0, 500.0, 3.0, 100.0, 0.0, 731.0])
    with torch.no_grad():
        out["fwd_z"], out["fwd_cond"], out["fwd_t"] = z.numpy(), cond.numpy(), t.numpy()
        out["fwd_out"] = head(z, t, cond).numpy()
    n = 3
    for algo in ("dpmsolver++", "sde-dpmsolver++"):
        for S in (1, 2, 5, 10, 20):
            c, ng, noise = synth.synth_inputs(n, cfg.hidden, cfg.latent, seed=S)
            torch.manual_seed(99)          # the SDE noise comes from the global CPU generator (dpm_solver.py:994-997)
            lat = ref_shim.ref_sample_speech_tokens(head, ref_shim.make_scheduler(algo), c, ng, 1.3, S, noise=noise)
            out[f"sample_{algo}_{S}"] = lat.numpy()
    # the reference in bf16 (model dtype of the GPU deployments): defines the tolerance budget
    head_bf = ref_head(cfg, W).to(torch.bfloat16)
    c, ng, noise = synth.synth_inputs(n, cfg.hidden, cfg.latent, seed=10)
    lat = ref_shim.ref_sample_speech_tokens(head_bf, ref_shim.make_scheduler(), c.bfloat16(), ng.bfloat16(), 1.3, 10,
                                            noise=noise)
    out["sample_bf16ref_dpmsolver++_10"] = lat.float().numpy()
    np.savez_compressed(os.path.join(HERE, "head_small.npz"), **out)

    # ---- decoder (structure-preserving tiny config, fp32) ----
    out = {}
    cfg = synth.DEC_TINY
    W = synth.init_decoder(cfg, seed=DEC_SEED)
    model = ref_acoustic(cfg, W)
    g = torch.Generator().manual_seed(1)
    for T in (1, 3, 7):
        lat = torch.randn(2, cfg.vae_dim, T, generator=g)
        with torch.no_grad():
            out[f"offline_lat_{T}"] = lat.numpy()
            out[f"offline_wave_{T}"] = model.decode(lat).numpy()
    ref = ref_shim.load()
    cache = ref.Cache()
    g = torch.Generator().manual_seed(2)
    k = 0
    for step in STREAM_PLAN:
        if step[0] == "zero":
            cache.set_to_zero(torch.tensor(step[1]))
            continue
        rows, T = step
        lat = torch.randn(len(rows), T, cfg.vae_dim, generator=g)
        with torch.no_grad():
            wave = model.decode(lat, cache=cache, sample_indices=torch.tensor(rows), use_cache=True)
        out[f"stream_lat_{k}"] = lat.numpy()
        out[f"stream_wave_{k}"] = wave.numpy()
        k += 1
    # reference-bf16 vs fp32 on the same input: the decoder's tolerance budget
    model_bf = ref_acoustic(cfg, W).to(torch.bfloat16)
    lat = torch.from_numpy(out["offline_lat_7"])
    with torch.no_grad():
        out["offline_wave_7_bf16ref"] = model_bf.decode(lat.bfloat16()).float().numpy()
    np.savez_compressed(os.path.join(HERE, "decoder_tiny.npz"), **out)
    for f in ("schedule.npz", "head_small.npz", "decoder_tiny.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
