#!/usr/bin/env python
"""Generates tests/golden/prefill_tiny.npz from the LIVE reference (/root/reference), CPU — the voice-prompt
prefill: acoustic_tokenizer.encode on prompts of ragged lengths, then
VibeVoiceTokenizerEncoderOutput.sample ('fix' and 'gaussian') and the bias/scale of _process_speech_inputs
(modeling_vibevoice_inference.py:150-164).

Run in the build container only:   python tests/golden/make_golden_prefill.py

Weights are re-derived by the tests from (seed, name) through vibevoice_b200/synth.py; stored are the seeded inputs,
the random draws the reference made (recorded by re-seeding torch's global generator) and the reference's OUTPUTS.
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

ENC_SEED = 7
LENGTHS = (1, 11, 12, 13, 50, 487)
BIAS, SCALE = -0.0703125, 0.1962890625          # exactly representable in bf16


def main():
    assert ref_shim.available(), "reference tree not found"
    out = {}
    cfg = synth.ACOUSTIC_ENC_TINY
    W = synth.init_encoder(cfg, seed=ENC_SEED)
    model = ref_shim.make_acoustic_encoder(cfg, W)
    for L in LENGTHS:
        audio = torch.randn(3, 1, L, generator=torch.Generator().manual_seed(L))
        with torch.no_grad():
            eo = model.encode(audio)
        out[f"audio_{L}"] = audio.numpy()
        out[f"mean_{L}"] = eo.mean.numpy()
    # the reference in bf16 on the longest prompt: the tolerance budget of a bf16 pipeline
    model_bf = ref_shim.make_acoustic_encoder(cfg, W).to(torch.bfloat16)
    with torch.no_grad():
        out["mean_487_bf16ref"] = model_bf.encode(torch.from_numpy(out["audio_487"]).bfloat16()).mean.float().numpy()
    # sampling + bias/scale, fp32 and bf16, on the mean of the 50-sample prompt
    mean32 = torch.from_numpy(out["mean_50"])
    b, s = torch.tensor(BIAS), torch.tensor(SCALE)
    for tag, mean in (("f32", mean32), ("bf16", mean32.bfloat16())):
        for dist in ("fix", "gaussian"):
            torch.manual_seed(21)
            x, std = ref_shim.encoder_output(mean, model.fix_std).sample(dist_type=dist)
            feats = (x + b) * s                                        # (:159)
            torch.manual_seed(21)
            if dist == "gaussian":
                out[f"row_noise_{dist}_{tag}"] = torch.randn(mean.shape[0], dtype=mean.dtype).float().numpy()
                out[f"std_{dist}_{tag}"] = std.float().numpy()
            out[f"noise_{dist}_{tag}"] = torch.randn_like(mean).float().numpy()
            out[f"feats_{dist}_{tag}"] = feats.float().numpy()
            assert feats.dtype == mean.dtype
    np.savez_compressed(os.path.join(HERE, "prefill_tiny.npz"), **out)
    print("wrote prefill_tiny.npz", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
