"""Host mirror of the acoustic tokenizer decoder and its streaming cache over libvvb200.

  * ``AcousticDecoderB200.decode(latents, cache=None, sample_indices=None, use_cache=False)``
        <- VibeVoiceAcousticTokenizerModel.decode (modular_vibevoice_tokenizer.py:1099-1108)
  * ``StreamingCacheB200`` with ``set_to_zero(sample_indices)`` / ``clear()``
        <- VibeVoiceTokenizerStreamingCache (modular_vibevoice_tokenizer.py:193-256)

Cache semantics: one HBM arena indexed by ``sample_idx`` (a row of the generation batch),
updated in place by the kernels; a slot that has never been stepped, or was zeroed, holds
zero history — exactly what the reference's "first chunk" path builds (:337-346, :488-492).
Rows are independent: a row joining late starts from zero history and does NOT reset the
others.  The reference's ``cache.get`` returns None if ANY row of the call is missing and its callers
then zero the history of EVERY row of the call (:204-207, :337-346, :488-492 — SURVEY.md §8(c) quirk
2C).  ``StreamingCacheB200(strict_reference_quirk=True)`` reproduces exactly that, for parity tests
only (it reads the indices on the host, one sync per call); the default is the sane behaviour.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import torch

from . import _lib, synth
from ._lib import check, current_stream, ptr


class StreamingCacheB200:
    """Device-arena replacement for VibeVoiceTokenizerStreamingCache (acoustic decoder layers).

    Keeps the reference's dict API (``cache`` attribute, ``get``/``set``) for any *other* module
    that shares the object (the semantic encoder stays on the torch path), and mirrors
    ``set_to_zero`` / ``clear`` into the arena.
    """

    def __init__(self, decoder: "AcousticDecoderB200" = None, max_slots: int = 64, strict_reference_quirk: bool = False):
        self.cache = {}  # dict API kept for non-accelerated layers
        self._dec = None
        self._c = None
        self.max_slots = int(max_slots)
        # TEST-ONLY: reproduce the reference's late-joiner behaviour (see the module docstring)
        self.strict_reference_quirk = bool(strict_reference_quirk)
        self._present = set()    # rows that have an entry in the reference's dict (strict mode only)
        if decoder is not None:
            self.bind(decoder)

    def bind(self, decoder: "AcousticDecoderB200") -> None:
        if self._c is not None:
            return
        self._dec = decoder
        c = C.c_void_p()
        with torch.cuda.device(decoder.device):
            check(decoder.lib.vv_cache_create(decoder._h, self.max_slots, C.byref(c)), "vv_cache_create")
        self._c = c

    def __del__(self):
        c = getattr(self, "_c", None)
        if c is not None and c.value and self._dec is not None:
            self._dec.lib.vv_cache_destroy(c)
            self._c = None

    # ---- reference API ----------------------------------------------------------------
    def get(self, layer_id, sample_indices):
        states = []
        for idx in sample_indices.tolist():
            key = (layer_id, idx)
            if key not in self.cache:
                return None
            states.append(self.cache[key])
        return torch.stack(states, dim=0)

    def set(self, layer_id, sample_indices, states):
        for i, idx in enumerate(sample_indices.tolist()):
            self.cache[(layer_id, idx)] = states[i].detach()

    def set_to_zero(self, sample_indices: torch.Tensor) -> None:
        """(:234-241) zero-fill the history of the given rows; no host sync for the arena part."""
        if len(self.cache):
            idx = set(sample_indices.tolist())
            for key in list(self.cache.keys()):
                if key[1] in idx:
                    self.cache[key] = torch.zeros_like(self.cache[key])
        if self._c is not None and sample_indices.numel() > 0:
            dev = self._dec.device
            si = sample_indices.to(dev, torch.int64).contiguous()
            with torch.cuda.device(dev):
                check(self._dec.lib.vv_cache_zero(self._c, ptr(si), si.numel(), current_stream()), "vv_cache_zero")

    def before_step(self, sample_indices: torch.Tensor) -> None:
        """Strict mode only: if any row of this call has no entry yet, every row of the call restarts from zero
        history, as in the reference (:204-207 -> :337-346 / :488-492); afterwards all of them have entries (:380, :547)."""
        if not self.strict_reference_quirk:
            return
        rows = [int(i) for i in sample_indices.tolist()]
        if any(r not in self._present for r in rows):
            self.set_to_zero(sample_indices)
        self._present.update(rows)

    def clear(self, layer_id=None, sample_indices=None) -> None:
        """(:243-256) dropping entries == zero history for the arena.  Only the full clear maps onto the arena: the
        reference's layer ids do not name arena regions, so a per-layer clear of a bound cache raises instead of
        silently keeping the old history (and `clear(None, sample_indices)` is a no-op in the reference too: none of its
        three branches matches)."""
        if layer_id is None and sample_indices is None:
            self.cache.clear()
            self._present.clear()
            if self._c is not None:
                with torch.cuda.device(self._dec.device):
                    check(self._dec.lib.vv_cache_zero_all(self._c, current_stream()), "vv_cache_zero_all")
        elif layer_id is not None:
            if self._c is not None and not any(k[0] == layer_id for k in self.cache):
                raise NotImplementedError(
                    "StreamingCacheB200.clear(layer_id=...): the device arena has no per-layer entries to drop; "
                    "use clear() or set_to_zero(sample_indices)")
            if sample_indices is None:
                for k in [k for k in self.cache if k[0] == layer_id]:
                    del self.cache[k]
            else:
                for idx in sample_indices.tolist():
                    self.cache.pop((layer_id, idx), None)


class AcousticDecoderB200:
    def __init__(self, weights: Dict[str, torch.Tensor], cfg: synth.DecCfg, device="cuda",
                 max_rows: int = 64, weight_prefix: str = ""):
        """weights: TokenizerDecoder state-dict (names relative to ``decoder.`` unless weight_prefix).
        max_rows: largest n*T (latent frames) one kernel pass handles; offline decodes longer than
        that are chunked in time inside the library."""
        self.lib = _lib.load()
        self.cfg = cfg
        self.device = torch.device(device)
        self.max_rows = int(max_rows)
        staged = [w.detach().to(self.device, torch.float32).contiguous()
                  for w in synth.ordered_weights(synth.dec_param_shapes(cfg), weights, weight_prefix, "decoder weight")]
        arr = (C.c_void_p * len(staged))(*[t.data_ptr() for t in staged])
        desc = _lib.VVDecDesc()
        desc.vae_dim = cfg.vae_dim
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
            check(self.lib.vv_dec_create(C.byref(desc), arr, len(staged), C.byref(h)), "vv_dec_create")
        self._h = h
        del staged

    @staticmethod
    def spec_from_module(acoustic_tokenizer):
        """(decoder state dict, DecCfg) of a live reference ``VibeVoiceAcousticTokenizerModel`` - host only."""
        dec = acoustic_tokenizer.decoder
        c = acoustic_tokenizer.config
        cfg = synth.DecCfg(vae_dim=c.vae_dim, n_filters=dec.n_filters, ratios=tuple(dec.ratios),
                           depths=tuple(dec.depths), kernel=7, eps=c.layernorm_eps)
        return dict(dec.state_dict()), cfg

    @classmethod
    def from_module(cls, acoustic_tokenizer, **kw):
        """Build from a live reference ``VibeVoiceAcousticTokenizerModel``."""
        W, cfg = cls.spec_from_module(acoustic_tokenizer)
        kw.setdefault("device", next(acoustic_tokenizer.decoder.parameters()).device)
        return cls(W, cfg, **kw)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            self.lib.vv_dec_destroy(h)
            self._h = None

    def new_cache(self, max_slots: int = 64, strict_reference_quirk: bool = False) -> StreamingCacheB200:
        return StreamingCacheB200(self, max_slots, strict_reference_quirk)

    @torch.no_grad()
    def decode(self, latents: torch.Tensor, cache: Optional[StreamingCacheB200] = None,
               sample_indices: Optional[torch.Tensor] = None, use_cache: bool = False, debug: bool = False,
               scale: float = 1.0, bias: float = 0.0) -> torch.Tensor:
        """latents [n,T,vae] or [n,vae,T] (permuted when shape[1] != vae_dim, :1102-1105) ->
        audio [n,1,hop*T] in the latents' dtype.  ``scale``/``bias`` fold the de-normalisation
        ``z / scale - bias`` of modeling_vibevoice_inference.py:651 into the stem."""
        if latents.dim() != 3:
            raise ValueError("latents must be [n, T, vae_dim] or [n, vae_dim, T]")
        V = self.cfg.vae_dim
        if latents.shape[1] == V:
            time_major, n, T = 0, latents.shape[0], latents.shape[2]
        else:
            if latents.shape[2] != V:
                raise ValueError(f"latents last dim {latents.shape[2]} != vae_dim {V}")
            time_major, n, T = 1, latents.shape[0], latents.shape[1]
        lat = latents.to(self.device)
        if lat.dtype == torch.bfloat16:
            io = _lib.VV_DTYPE_BF16
        else:
            lat = lat.to(torch.float32)
            io = _lib.VV_DTYPE_F32
        lat = lat.contiguous()
        wave = torch.empty(n, 1, self.cfg.hop * T, device=self.device, dtype=lat.dtype)
        with torch.cuda.device(self.device):
            if use_cache and cache is not None:
                # (:322-323, :473-474)
                assert sample_indices is not None, "sample_indices must be provided for streaming mode"
                assert len(sample_indices) == n, "sample_indices must match batch size"
                if cache._c is None:  # size the arena once, from the first batch seen (one sync, first frame only)
                    cache.max_slots = max(cache.max_slots, int(sample_indices.max()) + 1)
                cache.bind(self)
                if not sample_indices.is_cuda:
                    assert int(sample_indices.max()) < cache.max_slots, "sample index exceeds cache slots"
                si = sample_indices.to(self.device, torch.int64).contiguous()
                cache.before_step(sample_indices)
                check(self.lib.vv_dec_step(self._h, cache._c, ptr(lat), ptr(si), n, T, time_major,
                                           float(scale), float(bias), ptr(wave), io, current_stream()),
                      "vv_dec_step")
            else:
                check(self.lib.vv_dec_decode(self._h, ptr(lat), n, T, time_major, float(scale), float(bias),
                                             ptr(wave), io, current_stream()), "vv_dec_decode")
        return wave.to(latents.dtype)
