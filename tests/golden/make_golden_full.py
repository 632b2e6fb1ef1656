#!/usr/bin/env python
"""tests/golden/full_size.npz from the LIVE reference (/root/reference) at the REAL sizes of VibeVoice-1.5B, CPU:
one decoder row x 3 frames (offline decode and the same 3 frames streamed through the reference's cache) and one head
row through sample_speech_tokens with S = 10 — in fp32 (the truth) and in bf16 (the reference's own GPU precision: the
tolerance budget).  Run in the build container only (takes about two minutes):

    python tests/golden/make_golden_full.py

Weights are re-derived from (seed, name) by the tests (synth.py); only inputs and reference OUTPUTS are stored."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

from make_golden import ref_acoustic, ref_head  # noqa: E402
from oracle import ref_shim  # noqa: E402
from vibevoice_b200 import synth  # noqa: E402

SEED = 1234
FRAMES = 3


def main():
    assert ref_shim.available(), "reference tree not found"
    ref = ref_shim.load()
    out = {}
    # ---- decoder, full architecture, 1 row x 3 frames ----
    cfg = synth.DEC_FULL
    W = synth.init_decoder(cfg, seed=SEED)
    lat = torch.randn(1, FRAMES, cfg.vae_dim, generator=torch.Generator().manual_seed(41))
    out["dec_lat"] = lat.numpy()
    for dtype, tag in ((torch.float32, ""), (torch.bfloat16, "_bf16ref")):
        model = ref_acoustic(cfg, W).to(dtype)
        with torch.no_grad():
            out["dec_wave_offline" + tag] = model.decode(lat.to(dtype)).float().numpy()
            cache = ref.Cache()
            idx = torch.tensor([0])
            chunks = [model.decode(lat[:, f:f + 1].to(dtype), cache=cache, sample_indices=idx, use_cache=True) for f in range(FRAMES)]
            out["dec_wave_stream" + tag] = torch.cat(chunks, dim=-1).float().numpy()
        del model
    # ---- head, 1.5B configuration, 1 row (2 with CFG), S = 10, both solvers ----
    hcfg = synth.HEAD_1P5B
    Wh = synth.init_head(hcfg, seed=SEED)
    cond, neg, noise = synth.synth_inputs(1, hcfg.hidden, hcfg.latent, seed=SEED)
    for algo in ("dpmsolver++", "sde-dpmsolver++"):
        for dtype, tag in ((torch.float32, ""), (torch.bfloat16, "_bf16ref")):
            head = ref_head(hcfg, Wh).to(dtype)
            torch.manual_seed(99)      # the SDE noise comes from the global CPU generator (dpm_solver.py:994-997)
            latent = ref_shim.ref_sample_speech_tokens(head, ref_shim.make_scheduler(algo), cond.to(dtype), neg.to(dtype), 1.3, 10,
                                                       noise=noise)
            out[f"head_sample_{algo}{tag}"] = latent.float().numpy()
            del head
    np.savez_compressed(os.path.join(HERE, "full_size.npz"), **out)
    print({k: v.shape for k, v in out.items()}, os.path.getsize(os.path.join(HERE, "full_size.npz")), "bytes")


if __name__ == "__main__":
    main()
