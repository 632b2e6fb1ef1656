#!/usr/bin/env python
"""tests/golden/decoder_quirk_tiny.npz from the LIVE reference (/root/reference), CPU fp32: the late-joiner call
pattern through VibeVoiceAcousticTokenizerModel.decode with the reference's own
VibeVoiceTokenizerStreamingCache.  Run in the build container only:

    python tests/golden/make_golden_quirk.py

Only inputs-by-value and reference OUTPUTS are stored; weights are re-derived from (seed, name) by the tests."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

from make_golden import DEC_SEED, ref_acoustic  # noqa: E402
from oracle import ref_shim  # noqa: E402
from vibevoice_b200 import synth  # noqa: E402

# (rows, T) or ("zero", rows): rows 2 and 3 join a running stream
QUIRK_PLAN = [([0, 1], 1), ([0, 1], 1), ([0, 1, 2], 1), ([0, 1, 2], 1), ([1, 3], 2), ("zero", [0]), ([0, 1, 2, 3], 1)]


def main():
    assert ref_shim.available(), "reference tree not found"
    ref = ref_shim.load()
    cfg = synth.DEC_TINY
    W = synth.init_decoder(cfg, seed=DEC_SEED)
    out = {}
    for dtype, tag in ((torch.float32, ""), (torch.bfloat16, "_bf16ref")):
        model = ref_acoustic(cfg, W).to(dtype)
        cache = ref.Cache()
        g = torch.Generator().manual_seed(11)
        k = 0
        for step in QUIRK_PLAN:
            if step[0] == "zero":
                cache.set_to_zero(torch.tensor(step[1]))
                continue
            rows, T = step
            lat = torch.randn(len(rows), T, cfg.vae_dim, generator=g)
            with torch.no_grad():
                wave = model.decode(lat.to(dtype), cache=cache, sample_indices=torch.tensor(rows), use_cache=True)
            if tag == "":
                out[f"lat_{k}"] = lat.numpy()
            out[f"wave_{k}{tag}"] = wave.float().numpy()
            k += 1
    np.savez_compressed(os.path.join(HERE, "decoder_quirk_tiny.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
