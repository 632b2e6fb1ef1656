#!/usr/bin/env python
"""Generates tests/golden/encoder_tiny.npz and connector_small.npz from the LIVE reference (/root/reference),
CPU fp32 — the semantic-tokenizer encoder step and the SpeechConnector.

Run in the build container only:   python tests/golden/make_golden_encoder.py

Same rules as make_golden.py: weights are re-derived by the tests from (seed, name) through
vibevoice_b200/synth.py; only seeded inputs and the reference's OUTPUTS are stored.
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

ENC_SEED = 3
CONN_SEED = 2
CONN_SMALL = synth.ConnCfg(input_dim=16, output_dim=64)
STREAM_PLAN = [([0, 1, 2], 1), ([0, 1, 2], 1), ([0, 2], 1), ([0, 1, 2], 2), ("zero", [1]), ([0, 1, 2], 1), ([1], 1)]


def main():
    assert ref_shim.available(), "reference tree not found"
    ref = ref_shim.load()
    out = {}
    cfg = synth.ENC_TINY
    W = synth.init_encoder(cfg, seed=ENC_SEED)
    model = ref_shim.make_semantic(cfg, W)
    g = torch.Generator().manual_seed(1)
    for T in (1, 3, 6):
        audio = torch.randn(2, 1, cfg.hop * T, generator=g)
        with torch.no_grad():
            out[f"offline_audio_{T}"] = audio.numpy()
            out[f"offline_feat_{T}"] = model.encode(audio).mean.numpy()
    cache = ref.Cache()
    g = torch.Generator().manual_seed(2)
    k = 0
    for step in STREAM_PLAN:
        if step[0] == "zero":
            cache.set_to_zero(torch.tensor(step[1]))
            continue
        rows, T = step
        audio = torch.randn(len(rows), 1, cfg.hop * T, generator=g)
        with torch.no_grad():
            feat = model.encode(audio, cache=cache, sample_indices=torch.tensor(rows), use_cache=True).mean
        out[f"stream_audio_{k}"] = audio.numpy()
        out[f"stream_feat_{k}"] = feat.numpy()
        k += 1
    # the reference in bf16 on the same input: the tolerance budget
    model_bf = ref_shim.make_semantic(cfg, W).to(torch.bfloat16)
    audio = torch.from_numpy(out["offline_audio_6"])
    with torch.no_grad():
        out["offline_feat_6_bf16ref"] = model_bf.encode(audio.bfloat16()).mean.float().numpy()
    np.savez_compressed(os.path.join(HERE, "encoder_tiny.npz"), **out)

    out = {}
    W = synth.init_connector(CONN_SMALL, seed=CONN_SEED)
    m = ref_shim.make_connector(CONN_SMALL, W)
    x = torch.randn(5, 1, CONN_SMALL.input_dim, generator=torch.Generator().manual_seed(3))
    with torch.no_grad():
        out["x"] = x.numpy()
        out["y"] = m(x).numpy()
        out["y_bf16ref"] = ref_shim.make_connector(CONN_SMALL, W).to(torch.bfloat16)(x.bfloat16()).float().numpy()
    np.savez_compressed(os.path.join(HERE, "connector_small.npz"), **out)
    for f in ("encoder_tiny.npz", "connector_small.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
