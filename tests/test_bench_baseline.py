"""bench.py's `gpu_torch_baseline` leg (TorchGpuPath: the reference's eager torch path restated for timing on the GPU)
computes the same frame-step as the pinned oracle: tiny configuration, fp32, CPU, the same noise draw, three frames
through the streaming cache.  Keeps the baseline honest: it times the reference's arithmetic, not something cheaper."""
import importlib.util
import os

import torch

from oracle import vv_oracle as O
from vibevoice_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_torch_baseline_equals_oracle_tiny_fp32():
    bench = _bench()
    hcfg = synth.HeadCfg(hidden=64, latent=synth.DEC_TINY.vae_dim, layers=2, freq_dim=32)
    dcfg = synth.DEC_TINY
    n = 3
    path = bench.TorchGpuPath(n, torch.device("cpu"), hcfg=hcfg, dcfg=dcfg, dtype=torch.float32, seed=5)
    Wh, Wd = synth.init_head(hcfg, seed=5), synth.init_decoder(dcfg, seed=5)
    state = O.DecoderState(n)
    idx = torch.arange(n)
    for frame in range(3):
        torch.manual_seed(100 + frame)
        got = path.step()
        torch.manual_seed(100 + frame)
        noise = torch.randn(2 * n, hcfg.latent)                      # the draw step() makes (:716)
        lat = O.sample_speech_tokens(Wh, path.cond, path.neg, bench.CFG_SCALE, noise, steps=bench.DDPM_STEPS,
                                     t_dtype=torch.float32, layers=hcfg.layers, eps=hcfg.eps, freq_dim=hcfg.freq_dim)
        ref = O.decode(Wd, lat.unsqueeze(1), dcfg.depths, dcfg.ratios, dcfg.vae_dim, dcfg.eps, state=state, idx=idx)
        assert got.shape == ref.shape == (n, 1, dcfg.hop)
        assert ref.abs().max() > 1e-3
        torch.testing.assert_close(got, ref, rtol=2e-3, atol=2e-4 * ref.abs().max().item())
