"""Architecture descriptions + deterministic "observable" random init for the hot path.

There is no network for checkpoints, so tests, bench.py and smoke() use random-init weights
of the named architecture.  The reference's default init hides bugs (adaLN / final linear are
zero — modular_vibevoice_diffusion_head.py:246-252 — and the decoder's layer-scale is 1e-6 —
modular_vibevoice_tokenizer.py:658-660), so every parameter is re-drawn here ("observable
init").  Each tensor has its own generator seeded from (seed, crc32(name)):
the values depend only on the parameter's name and shape, not on construction order, so the
golden fixtures, the oracle, and the CUDA path all see identical weights.
"""
from __future__ import annotations

import zlib
from collections import OrderedDict
from dataclasses import dataclass
from typing import Dict, Tuple

import torch


@dataclass(frozen=True)
class HeadCfg:
    """VibeVoiceDiffusionHeadConfig subset (configuration_vibevoice.py:130-162)."""
    hidden: int = 1536
    ffn_ratio: float = 3.0
    latent: int = 64
    layers: int = 4
    freq_dim: int = 256
    eps: float = 1e-5

    @property
    def ffn(self) -> int:
        return int(self.hidden * self.ffn_ratio)


HEAD_1P5B = HeadCfg()
HEAD_LARGE = HeadCfg(hidden=3584)


@dataclass(frozen=True)
class DecCfg:
    """Decoder half of VibeVoiceAcousticTokenizerConfig (configuration_vibevoice.py:13-74)."""
    vae_dim: int = 64
    n_filters: int = 32
    ratios: Tuple[int, ...] = (8, 5, 5, 4, 2, 2)
    depths: Tuple[int, ...] = (8, 3, 3, 3, 3, 3, 3)
    kernel: int = 7
    eps: float = 1e-5

    @property
    def n_stages(self) -> int:
        return len(self.depths)

    def channels(self, i: int) -> int:
        return self.n_filters * 2 ** (self.n_stages - 1 - i)

    @property
    def hop(self) -> int:
        h = 1
        for r in self.ratios:
            h *= r
        return h


DEC_FULL = DecCfg()
# structure-preserving small decoder for fast CPU tests (hop = 2*2*2 = 8)
DEC_TINY = DecCfg(vae_dim=16, n_filters=8, ratios=(2, 2, 2), depths=(2, 1, 1, 2))


@dataclass(frozen=True)
class EncCfg:
    """Encoder of VibeVoiceSemanticTokenizerConfig (configuration_vibevoice.py:76-127).  ``ratios`` is in the
    order TokenizerEncoder applies them (it reverses config.encoder_ratios, modular_vibevoice_tokenizer.py:701)."""
    dim: int = 128
    n_filters: int = 32
    ratios: Tuple[int, ...] = (2, 2, 4, 5, 5, 8)
    depths: Tuple[int, ...] = (3, 3, 3, 3, 3, 3, 8)
    kernel: int = 7
    eps: float = 1e-5

    @property
    def n_stages(self) -> int:
        return len(self.depths)

    def channels(self, i: int) -> int:
        return self.n_filters * 2 ** i

    @property
    def hop(self) -> int:
        h = 1
        for r in self.ratios:
            h *= r
        return h


ENC_FULL = EncCfg()
# structure-preserving small encoder for fast CPU tests (hop = 8)
ENC_TINY = EncCfg(dim=16, n_filters=8, ratios=(2, 2, 2), depths=(2, 1, 1, 2))
# Acoustic-tokenizer encoder (VibeVoiceAcousticTokenizerConfig, configuration_vibevoice.py:13-73): the same stack with
# vae_dim 64; its output is the mean of a fixed-variance Gaussian (fix_std 0.5, std_dist_type 'gaussian').
ACOUSTIC_ENC_FULL = EncCfg(dim=64)
# small prompt encoder with unequal strides (hop = 12) so that ragged prompt lengths exercise every stage's padding
ACOUSTIC_ENC_TINY = EncCfg(dim=16, n_filters=8, ratios=(2, 3, 2), depths=(2, 1, 1, 2))


@dataclass(frozen=True)
class ConnCfg:
    """SpeechConnector (modeling_vibevoice.py:58-69): Linear(in, out) -> LlamaRMSNorm(out, 1e-6) -> Linear(out, out)."""
    input_dim: int = 64
    output_dim: int = 1536
    eps: float = 1e-6


def enc_param_shapes(cfg: EncCfg) -> "OrderedDict[str, Tuple[int, ...]]":
    """TokenizerEncoder state-dict names in the order vv_enc_create expects (include/vvb200.h)."""
    k = cfg.kernel
    d: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    d["downsample_layers.0.0.conv.conv.weight"] = (cfg.channels(0), 1, k)
    d["downsample_layers.0.0.conv.conv.bias"] = (cfg.channels(0),)
    for i in range(cfg.n_stages):
        c = cfg.channels(i)
        if i > 0:
            r = cfg.ratios[i - 1]
            d[f"downsample_layers.{i}.0.conv.conv.weight"] = (c, cfg.channels(i - 1), 2 * r)
            d[f"downsample_layers.{i}.0.conv.conv.bias"] = (c,)
        for j in range(cfg.depths[i]):
            p = f"stages.{i}.{j}."
            d[p + "norm.weight"] = (c,)
            d[p + "mixer.conv.conv.conv.weight"] = (c, 1, k)
            d[p + "mixer.conv.conv.conv.bias"] = (c,)
            d[p + "gamma"] = (c,)
            d[p + "ffn_norm.weight"] = (c,)
            d[p + "ffn.linear1.weight"] = (4 * c, c)
            d[p + "ffn.linear1.bias"] = (4 * c,)
            d[p + "ffn.linear2.weight"] = (c, 4 * c)
            d[p + "ffn.linear2.bias"] = (c,)
            d[p + "ffn_gamma"] = (c,)
    cl = cfg.channels(cfg.n_stages - 1)
    d["head.conv.conv.weight"] = (cfg.dim, cl, k)
    d["head.conv.conv.bias"] = (cfg.dim,)
    return d


def conn_param_shapes(cfg: ConnCfg) -> "OrderedDict[str, Tuple[int, ...]]":
    d: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    d["fc1.weight"] = (cfg.output_dim, cfg.input_dim)
    d["fc1.bias"] = (cfg.output_dim,)
    d["norm.weight"] = (cfg.output_dim,)
    d["fc2.weight"] = (cfg.output_dim, cfg.output_dim)
    d["fc2.bias"] = (cfg.output_dim,)
    return d


def head_param_shapes(cfg: HeadCfg) -> "OrderedDict[str, Tuple[int, ...]]":
    """State-dict names in the order vv_head_create expects (include/vvb200.h)."""
    H, F, L = cfg.hidden, cfg.ffn, cfg.latent
    d: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    d["noisy_images_proj.weight"] = (H, L)
    d["cond_proj.weight"] = (H, H)
    d["t_embedder.mlp.0.weight"] = (H, cfg.freq_dim)
    d["t_embedder.mlp.2.weight"] = (H, H)
    for i in range(cfg.layers):
        d[f"layers.{i}.norm.weight"] = (H,)
        d[f"layers.{i}.adaLN_modulation.1.weight"] = (3 * H, H)
        d[f"layers.{i}.ffn.gate_proj.weight"] = (F, H)
        d[f"layers.{i}.ffn.up_proj.weight"] = (F, H)
        d[f"layers.{i}.ffn.down_proj.weight"] = (H, F)
    d["final_layer.adaLN_modulation.1.weight"] = (2 * H, H)
    d["final_layer.linear.weight"] = (L, H)
    return d


def dec_param_shapes(cfg: DecCfg) -> "OrderedDict[str, Tuple[int, ...]]":
    """TokenizerDecoder state-dict names (relative to ``decoder.``) in vv_dec_create order."""
    k = cfg.kernel
    d: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    c0 = cfg.channels(0)
    d["upsample_layers.0.0.conv.conv.weight"] = (c0, cfg.vae_dim, k)
    d["upsample_layers.0.0.conv.conv.bias"] = (c0,)
    for i in range(cfg.n_stages):
        c = cfg.channels(i)
        if i > 0:
            r = cfg.ratios[i - 1]
            d[f"upsample_layers.{i}.0.convtr.convtr.weight"] = (cfg.channels(i - 1), c, 2 * r)
            d[f"upsample_layers.{i}.0.convtr.convtr.bias"] = (c,)
        for j in range(cfg.depths[i]):
            p = f"stages.{i}.{j}."
            d[p + "norm.weight"] = (c,)
            d[p + "mixer.conv.conv.conv.weight"] = (c, 1, k)
            d[p + "mixer.conv.conv.conv.bias"] = (c,)
            d[p + "gamma"] = (c,)
            d[p + "ffn_norm.weight"] = (c,)
            d[p + "ffn.linear1.weight"] = (4 * c, c)
            d[p + "ffn.linear1.bias"] = (4 * c,)
            d[p + "ffn.linear2.weight"] = (c, 4 * c)
            d[p + "ffn.linear2.bias"] = (c,)
            d[p + "ffn_gamma"] = (c,)
    cl = cfg.channels(cfg.n_stages - 1)
    d["head.conv.conv.weight"] = (1, cl, k)
    d["head.conv.conv.bias"] = (1,)
    return d


def _gen(seed: int, name: str) -> torch.Generator:
    return torch.Generator().manual_seed((seed * 1000003 + zlib.crc32(name.encode())) & 0x7FFFFFFF)


def init_head(cfg: HeadCfg, seed: int = 1234) -> Dict[str, torch.Tensor]:
    """fp32 CPU weights: matrices ~ N(0, 0.02^2) (incl. the zero-init ones), norms 1+0.1N."""
    out = OrderedDict()
    for name, shape in head_param_shapes(cfg).items():
        g = _gen(seed, name)
        if len(shape) == 1:
            out[name] = 1.0 + 0.1 * torch.randn(shape, generator=g)
        else:
            out[name] = 0.02 * torch.randn(shape, generator=g)
    return out


def init_decoder(cfg: DecCfg, seed: int = 1234) -> Dict[str, torch.Tensor]:
    """fp32 CPU weights: conv/linear ~ N(0, 1/fan_in), bias ~ N(0, 0.02^2), layer-scale
    U(0.2, 0.6), norm weights 1 + 0.1 N(0, 1)."""
    out = OrderedDict()
    for name, shape in dec_param_shapes(cfg).items():
        g = _gen(seed, name)
        if name.endswith("gamma"):
            out[name] = 0.2 + 0.4 * torch.rand(shape, generator=g)
        elif name.endswith("norm.weight"):
            out[name] = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif name.endswith(".bias"):
            out[name] = 0.02 * torch.randn(shape, generator=g)
        elif "convtr" in name:
            cin, _, kk = shape
            fan_in = 2 * cin  # k = 2*stride: every output sample sums 2 taps x Cin
            out[name] = torch.randn(shape, generator=g) / fan_in ** 0.5
        elif len(shape) == 3:
            fan_in = shape[1] * shape[2]
            out[name] = torch.randn(shape, generator=g) / fan_in ** 0.5
        else:
            out[name] = torch.randn(shape, generator=g) / shape[1] ** 0.5
    return out


def init_encoder(cfg: EncCfg, seed: int = 1234) -> Dict[str, torch.Tensor]:
    """fp32 CPU weights, same recipe as init_decoder (names are prefixed so the two tokenizers differ)."""
    out = OrderedDict()
    for name, shape in enc_param_shapes(cfg).items():
        g = _gen(seed, "semantic_encoder." + name)
        if name.endswith("gamma"):
            out[name] = 0.2 + 0.4 * torch.rand(shape, generator=g)
        elif name.endswith("norm.weight"):
            out[name] = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif name.endswith(".bias"):
            out[name] = 0.02 * torch.randn(shape, generator=g)
        elif len(shape) == 3:
            out[name] = torch.randn(shape, generator=g) / (shape[1] * shape[2]) ** 0.5
        else:
            out[name] = torch.randn(shape, generator=g) / shape[1] ** 0.5
    return out


def init_connector(cfg: ConnCfg, seed: int = 1234, tag: str = "acoustic") -> Dict[str, torch.Tensor]:
    out = OrderedDict()
    for name, shape in conn_param_shapes(cfg).items():
        g = _gen(seed, f"connector.{tag}." + name)
        if name == "norm.weight":
            out[name] = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif name.endswith(".bias"):
            out[name] = 0.02 * torch.randn(shape, generator=g)
        else:
            out[name] = torch.randn(shape, generator=g) / shape[1] ** 0.5
    return out


def ordered_weights(shapes, weights, prefix: str = "", what: str = "weight"):
    """The tensors of ``weights`` in the order of ``shapes`` (an *_param_shapes dict), checked for presence and shape:
    everything a handle constructor does with a state dict before the device copy (no CUDA needed)."""
    out = []
    for name, shape in shapes.items():
        key = prefix + name
        if key not in weights:
            raise KeyError(f"missing {what} {key}")
        w = weights[key]
        if tuple(w.shape) != tuple(shape):
            raise ValueError(f"{key}: expected shape {tuple(shape)}, got {tuple(w.shape)}")
        out.append(w)
    return out


def synth_inputs(n: int, hidden: int, latent: int, seed: int = 1234):
    """cond, neg ~ N(0,1) [n,H]; noise ~ N(0,1) [2n,L] drawn like modeling_vibevoice_inference.py:716."""
    g = _gen(seed, f"inputs.{n}.{hidden}.{latent}")
    cond = torch.randn(n, hidden, generator=g)
    neg = torch.randn(n, hidden, generator=g)
    noise = torch.randn(2 * n, latent, generator=g)
    return cond, neg, noise
