"""Host mirror of the tokenizer encoders and the SpeechConnector over libvvb200
(the two ops that follow ``acoustic_tokenizer.decode`` inside one frame
iteration of ``generate()``, modeling_vibevoice_inference.py:673-687; rank 4: the voice-prompt prefill,
``_process_speech_inputs`` :150-164).

  * ``SemanticEncoderB200.encode(audio, cache=None, sample_indices=None, use_cache=False).mean``
        <- VibeVoiceSemanticTokenizerModel.encode (modular_vibevoice_tokenizer.py:1167-1170)
  * ``SpeechConnectorB200(features)``  <- SpeechConnector.forward (modeling_vibevoice.py:58-69)
  * ``AcousticEncoderB200.encode(audio)`` -> ``EncoderOutputB200(mean, std)`` with ``sample('fix'|'gaussian')``
        <- VibeVoiceAcousticTokenizerModel.encode / VibeVoiceTokenizerEncoderOutput.sample (:1081-1085, :954-1000)

The cache is a ``StreamingCacheB200`` bound to the encoder (one arena per tokenizer, like the reference's
separate ``semantic_cache`` object, modeling_vibevoice_inference.py:388).  With a cache the audio length must be a
whole number of frames (hop = 3200 samples): that is what the frame loop feeds.  Without one any length is taken and
the last frame is completed the way the reference's right "extra padding" does (:127-133).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import torch

from . import _lib, synth
from ._lib import check, current_stream, ptr
from .decoder import StreamingCacheB200



def encoder_spec_from_module(tokenizer):
    """(encoder state dict, EncCfg) of a live reference semantic / acoustic tokenizer model - host only, no CUDA."""
    enc = tokenizer.encoder
    c = tokenizer.config
    cfg = synth.EncCfg(dim=c.vae_dim, n_filters=enc.n_filters, ratios=tuple(enc.ratios), depths=tuple(enc.depths), kernel=7,
                       eps=c.layernorm_eps)
    return dict(enc.state_dict()), cfg


def connector_spec_from_module(connector):
    """(state dict, ConnCfg) of a live reference ``SpeechConnector`` - host only, no CUDA."""
    cfg = synth.ConnCfg(input_dim=connector.fc1.in_features, output_dim=connector.fc1.out_features,
                        eps=connector.norm.variance_epsilon)
    return dict(connector.state_dict()), cfg


class EncoderOutputB200:
    """Stand-in for VibeVoiceTokenizerEncoderOutput (modular_vibevoice_tokenizer.py:954-1000): ``mean`` [n, T, dim]
    and the fixed ``std`` (None for the semantic tokenizer, whose std_dist_type is 'none', :1172-1175).
    ``sample(dist_type)`` draws its noise from torch's generator of the mean's device in the reference's order
    (:980-990: 'gaussian' draws randn(batch) and then randn_like(mean); 'fix' only the latter), so a seeded run sees
    the same random numbers, and does the arithmetic in one kernel (``vv_latent_sample``)."""

    def __init__(self, mean: torch.Tensor, std=None):
        self.mean = mean
        self.std = std

    def _sample(self, dist_type: str, bias: float, scale: float):
        mean = self.mean
        if not mean.is_cuda:
            raise RuntimeError("EncoderOutputB200.sample needs the mean on a CUDA device")
        lib = _lib.load()
        m = mean if mean.dtype in (torch.bfloat16, torch.float32) else mean.to(torch.float32)
        m = m.contiguous()
        io = _lib.VV_DTYPE_BF16 if m.dtype == torch.bfloat16 else _lib.VV_DTYPE_F32
        n = m.shape[0]
        per_row = m[0].numel()
        noise = row_noise = std_out = None
        std_arg = 0.0
        used = self.std
        if dist_type == "fix" and self.std is not None:
            noise = torch.randn_like(m)
            std_arg = float(self.std)
        elif dist_type == "gaussian" and self.std is not None:
            row_noise = torch.randn(n, device=m.device, dtype=m.dtype)
            noise = torch.randn_like(m)
            std_arg = float(self.std) / 0.8
            std_out = torch.empty(n, device=m.device, dtype=m.dtype)
        out = torch.empty_like(m)
        with torch.cuda.device(m.device):
            check(lib.vv_latent_sample(ptr(m), ptr(noise) if noise is not None else None,
                                       ptr(row_noise) if row_noise is not None else None, std_arg, float(bias),
                                       float(scale), n, 1, per_row, io, ptr(out),
                                       ptr(std_out) if std_out is not None else None, current_stream()),
                  "vv_latent_sample")
        if std_out is not None:
            used = std_out.reshape([n] + [1] * (m.dim() - 1))
        return out.to(mean.dtype), used

    def sample(self, dist_type: str = "fix"):
        """(:970-994) -> (x, std used).  Any dist_type other than 'fix' / 'gaussian' returns the mean."""
        if dist_type not in ("fix", "gaussian") or self.std is None:
            return self.mean, self.std
        return self._sample(dist_type, 0.0, 1.0)

    def sample_features(self, dist_type: str, bias_factor, scaling_factor):
        """``(sample(dist_type)[0] + speech_bias_factor) * speech_scaling_factor`` in the same launch
        (modeling_vibevoice_inference.py:155-159)."""
        return self._sample(dist_type, float(bias_factor), float(scaling_factor))[0]

    def mode(self):
        return self.mean


class SemanticEncoderB200:
    fix_std = None            # the semantic tokenizer has no VAE head (fix_std 0, std_dist_type 'none')
    std_dist_type = "none"

    def __init__(self, weights: Dict[str, torch.Tensor], cfg: synth.EncCfg, device="cuda", max_rows: int = 64,
                 weight_prefix: str = ""):
        """weights: TokenizerEncoder state-dict (names relative to ``encoder.`` unless weight_prefix).
        max_rows: largest n*T (frames) one call handles."""
        self.lib = _lib.load()
        self.cfg = cfg
        self.device = torch.device(device)
        self.max_rows = int(max_rows)
        staged = [w.detach().to(self.device, torch.float32).contiguous()
                  for w in synth.ordered_weights(synth.enc_param_shapes(cfg), weights, weight_prefix, "encoder weight")]
        arr = (C.c_void_p * len(staged))(*[t.data_ptr() for t in staged])
        desc = _lib.VVDecDesc()
        desc.vae_dim = cfg.dim
        desc.n_stages = cfg.n_stages
        for i in range(cfg.n_stages):
            desc.channels[i] = cfg.channels(i)
            desc.depths[i] = cfg.depths[i]
            desc.ratios[i] = cfg.ratios[i - 1] if i > 0 else 1
        desc.kernel = cfg.kernel
        desc.eps = cfg.eps
        desc.weight_dtype = _lib.VV_DTYPE_F32
        desc.max_rows = self.max_rows
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            torch.cuda.synchronize()
            check(self.lib.vv_enc_create(C.byref(desc), arr, len(staged), C.byref(h)), "vv_enc_create")
        self._h = h
        del staged

    @classmethod
    def from_module(cls, semantic_tokenizer, **kw):
        """Build from a live reference ``VibeVoiceSemanticTokenizerModel``."""
        W, cfg = encoder_spec_from_module(semantic_tokenizer)
        kw.setdefault("device", next(semantic_tokenizer.encoder.parameters()).device)
        return cls(W, cfg, **kw)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            self.lib.vv_enc_destroy(h)
            self._h = None

    def new_cache(self, max_slots: int = 64) -> StreamingCacheB200:
        return StreamingCacheB200(self, max_slots)

    @torch.no_grad()
    def encode(self, audio: torch.Tensor, cache: Optional[StreamingCacheB200] = None,
               sample_indices: Optional[torch.Tensor] = None, use_cache: bool = False,
               debug: bool = False) -> EncoderOutputB200:
        """audio [n, 1, hop*T] (or [n, hop*T]) -> ``.mean`` [n, T, dim] in the audio's dtype.  Without a cache the
        length may be anything: the last frame is completed with zeros stage by stage (``vv_enc_prefill``)."""
        if audio.dim() == 3:
            if audio.shape[1] != 1:
                raise ValueError("audio must be [n, 1, samples]")
            n, L = audio.shape[0], audio.shape[2]
        elif audio.dim() == 2:
            n, L = audio.shape
        else:
            raise ValueError("audio must be [n, 1, samples] or [n, samples]")
        hop = self.cfg.hop
        streaming = use_cache and cache is not None
        if L == 0 or (streaming and L % hop != 0):
            raise ValueError(f"audio length {L} is not a whole number of frames (hop {hop})")
        T = -(-L // hop)
        a = audio.to(self.device)
        if a.dtype == torch.bfloat16:
            io = _lib.VV_DTYPE_BF16
        else:
            a = a.to(torch.float32)
            io = _lib.VV_DTYPE_F32
        a = a.contiguous()
        out = torch.empty(n, T, self.cfg.dim, device=self.device, dtype=a.dtype)
        with torch.cuda.device(self.device):
            if streaming:
                # (:322-323)
                assert sample_indices is not None, "sample_indices must be provided for streaming mode"
                assert len(sample_indices) == n, "sample_indices must match batch size"
                if cache._c is None:
                    cache.max_slots = max(cache.max_slots, int(sample_indices.max()) + 1)
                cache.bind(self)
                if not sample_indices.is_cuda:
                    assert int(sample_indices.max()) < cache.max_slots, "sample index exceeds cache slots"
                si = sample_indices.to(self.device, torch.int64).contiguous()
                if hasattr(cache, "before_step"):
                    cache.before_step(sample_indices)
                check(self.lib.vv_enc_step(self._h, cache._c, ptr(a), ptr(si), n, T, ptr(out), io, current_stream()),
                      "vv_enc_step")
            else:
                if n > self.max_rows:
                    raise ValueError(f"batch {n} exceeds max_rows {self.max_rows}")
                check(self.lib.vv_enc_prefill(self._h, ptr(a), n, L, ptr(out), io, current_stream()), "vv_enc_prefill")
        return EncoderOutputB200(out.to(audio.dtype), self.fix_std)


class AcousticEncoderB200(SemanticEncoderB200):
    """The acoustic tokenizer's encoder for the voice-prompt prefill:
    ``encode(audio)`` <- VibeVoiceAcousticTokenizerModel.encode (modular_vibevoice_tokenizer.py:1081-1085), whose
    output carries ``std = fix_std``; ``sampling`` <- (:1087-1097)."""

    def __init__(self, weights, cfg: synth.EncCfg, fix_std: float = 0.5, std_dist_type: str = "gaussian", **kw):
        super().__init__(weights, cfg, **kw)
        self.fix_std = fix_std
        self.std_dist_type = std_dist_type

    @classmethod
    def from_module(cls, acoustic_tokenizer, **kw):
        """Build from a live reference ``VibeVoiceAcousticTokenizerModel``."""
        W, cfg = encoder_spec_from_module(acoustic_tokenizer)
        kw.setdefault("device", next(acoustic_tokenizer.encoder.parameters()).device)
        return cls(W, cfg, fix_std=acoustic_tokenizer.fix_std, std_dist_type=acoustic_tokenizer.std_dist_type, **kw)

    def sampling(self, encoder_output: EncoderOutputB200, dist_type: Optional[str] = None):
        dist_type = dist_type or self.std_dist_type
        if dist_type not in ("fix", "gaussian"):
            raise ValueError(f"Unsupported dist_type: {dist_type}, expected 'fix' or 'gaussian'")
        return encoder_output.sample(dist_type=dist_type)


class SpeechConnectorB200:
    def __init__(self, weights: Dict[str, torch.Tensor], cfg: synth.ConnCfg, device="cuda", max_rows: int = 64):
        self.lib = _lib.load()
        self.cfg = cfg
        self.device = torch.device(device)
        self.max_rows = int(max_rows)
        staged = [w.detach().to(self.device, torch.float32).contiguous()
                  for w in synth.ordered_weights(synth.conn_param_shapes(cfg), weights, what="connector weight")]
        arr = (C.c_void_p * len(staged))(*[t.data_ptr() for t in staged])
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            torch.cuda.synchronize()
            check(self.lib.vv_conn_create(cfg.input_dim, cfg.output_dim, float(cfg.eps), self.max_rows,
                                          _lib.VV_DTYPE_F32, arr, len(staged), C.byref(h)), "vv_conn_create")
        self._h = h
        del staged

    @classmethod
    def from_module(cls, connector, **kw):
        """Build from a live reference ``SpeechConnector``."""
        W, cfg = connector_spec_from_module(connector)
        kw.setdefault("device", connector.fc1.weight.device)
        return cls(W, cfg, **kw)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            self.lib.vv_conn_destroy(h)
            self._h = None

    @torch.no_grad()
    def __call__(self, features: torch.Tensor, addend: Optional[torch.Tensor] = None, **kwargs) -> torch.Tensor:
        """features [..., in_dim] -> [..., out_dim] (bf16 on the device; cast back to the input dtype).
        ``addend`` (same shape as the result) is summed in the last epilogue: acoustic_embed + semantic_embed."""
        lead = features.shape[:-1]
        x = features.reshape(-1, features.shape[-1]).to(self.device)
        if x.shape[1] != self.cfg.input_dim:
            raise ValueError(f"features last dim {x.shape[1]} != input_dim {self.cfg.input_dim}")
        if x.dtype == torch.bfloat16:
            io = _lib.VV_DTYPE_BF16
        else:
            x = x.to(torch.float32)
            io = _lib.VV_DTYPE_F32
        x = x.contiguous()
        M = x.shape[0]
        out = torch.empty(M, self.cfg.output_dim, device=self.device, dtype=torch.bfloat16)
        add = None
        if addend is not None:
            add = addend.reshape(M, self.cfg.output_dim).to(self.device, torch.bfloat16).contiguous()
        with torch.cuda.device(self.device):
            for m0 in range(0, M, self.max_rows):        # a voice prompt is n*T rows: chunk by the handle's capacity
                m1 = min(M, m0 + self.max_rows)
                check(self.lib.vv_conn_forward(self._h, ptr(x[m0:m1]), m1 - m0, io,
                                               ptr(add[m0:m1]) if add is not None else None, ptr(out[m0:m1]),
                                               current_stream()), "vv_conn_forward")
        return out.reshape(*lead, self.cfg.output_dim).to(features.dtype)

    forward = __call__
