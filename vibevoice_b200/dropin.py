"""Drop-in installation onto a live reference model.

    import vibevoice_b200.dropin as vvb
    handle = vvb.install(model)        # model: VibeVoiceForConditionalGenerationInference on cuda
    outputs = model.generate(**inputs, cfg_scale=1.3, tokenizer=processor.tokenizer)   # unchanged
    handle.uninstall()

``generate()`` (modeling_vibevoice_inference.py:327-710) is untouched; the three callables it
reaches on the acoustic path are rebound on the INSTANCE, with the reference's signatures:

  * ``model.sample_speech_tokens(condition, neg_condition, cfg_scale=3.0)``            (:712-725)
  * ``model.model.prediction_head.forward(noisy_images, timesteps, condition)``         (diffusion_head.py:254-259)
  * ``model.model.acoustic_tokenizer.decode(latents, cache=None, sample_indices=None,
        use_cache=False, debug=False)``                                                (tokenizer.py:1099-1108)

and the cache object ``generate()`` constructs (``VibeVoiceTokenizerStreamingCache()``, :387) gets
a device arena attached on first use; its ``set_to_zero`` (:559) is mirrored into the arena.  The
semantic encoder keeps using the same object's dict API.  The LLM and the streamer stay on the reference
torch path.  There is no fallback: if libvvb200.so is missing ``install`` raises.

``install(model, frame_tail=True)`` additionally rebinds the rest of the frame iteration (SURVEY.md §8(f)
ranks 1-2, modeling_vibevoice_inference.py:673-687):

  * ``model.model.semantic_tokenizer.encode(audio, cache=None, sample_indices=None, use_cache=False,
        debug=False)`` -> object with ``.mean``                                       (tokenizer.py:1167-1170)
  * ``model.model.acoustic_connector(features)`` / ``model.model.semantic_connector(features)``
                                                                                       (modeling_vibevoice.py:58-69)

The semantic cache object ``generate()`` constructs (:388) gets its own arena, bound to the encoder.

``install(model, prefill=True)`` rebinds the voice-prompt path:

  * ``model.model.acoustic_tokenizer.encode(audio)`` (no cache) -> object with ``.mean``, ``.std``, ``.sample``
                                                                                       (tokenizer.py:1081-1085)
  * ``model._process_speech_inputs(speech_tensors, speech_masks, speech_type="audio")``
        -> (acoustic_features, acoustic_connected)                        (modeling_vibevoice_inference.py:150-177)
"""
from __future__ import annotations

import types
import weakref
from typing import Optional

import torch

from . import _lib
from .decoder import AcousticDecoderB200, StreamingCacheB200
from .encoder import AcousticEncoderB200, EncoderOutputB200, SemanticEncoderB200, SpeechConnectorB200
from .head import DiffusionHeadB200


class Installed:
    def __init__(self, model, head: DiffusionHeadB200, dec: AcousticDecoderB200, max_batch: int):
        self.model, self.head, self.dec, self.max_batch = model, head, dec, max_batch
        self._saved = []                              # (object, attribute, previous instance value)
        self._arenas = weakref.WeakKeyDictionary()   # reference cache object -> StreamingCacheB200

    # ---- cache adoption ---------------------------------------------------------------------
    def arena_for(self, cache, owner=None) -> StreamingCacheB200:
        """Our own cache class is used as is; a reference VibeVoiceTokenizerStreamingCache gets an
        arena attached and its set_to_zero / clear mirrored into it.  ``owner``: the tokenizer handle the
        arena belongs to (default: the acoustic decoder)."""
        owner = owner if owner is not None else self.dec
        if isinstance(cache, StreamingCacheB200):
            cache.bind(owner)
            return cache
        arena = self._arenas.get(cache)
        if arena is None:
            arena = StreamingCacheB200(owner, max_slots=self.max_batch)
            self._arenas[cache] = arena
            ref_zero, ref_clear = cache.set_to_zero, getattr(cache, "clear", None)

            def set_to_zero(_self, sample_indices, _ref=ref_zero, _arena=arena):
                _ref(sample_indices)                 # dict entries of the other (torch-path) layers
                _arena.set_to_zero(sample_indices)

            cache.set_to_zero = types.MethodType(set_to_zero, cache)
            if ref_clear is not None:
                def clear(_self, layer_id=None, sample_indices=None, _ref=ref_clear, _arena=arena):
                    _ref(layer_id, sample_indices)
                    if layer_id is None and sample_indices is None:
                        _arena.clear()

                cache.clear = types.MethodType(clear, cache)
        return arena

    def uninstall(self) -> None:
        for obj, name, val in reversed(self._saved):
            if val is None:
                try:
                    delattr(obj, name)
                except AttributeError:
                    pass
            else:
                setattr(obj, name, val)
        self._saved.clear()

    def _bind(self, obj, name, fn) -> None:
        self._saved.append((obj, name, obj.__dict__.get(name)))   # None => class attribute underneath
        setattr(obj, name, fn)


def _algorithm_type(scheduler) -> str:
    cfg = getattr(scheduler, "config", None)
    algo = getattr(cfg, "algorithm_type", None) if cfg is not None else None
    if algo is None and isinstance(cfg, dict):
        algo = cfg.get("algorithm_type")
    return algo or "dpmsolver++"


# what schedule.py / vv_head_set_schedule implement (modeling_vibevoice.py:138-142 + the defaults of
# DPMSolverMultistepScheduler, schedule/dpm_solver.py:203-227); anything else must fail loudly, not drift
_SCHEDULER_EXPECTED = {"prediction_type": ("v_prediction",), "beta_schedule": ("cosine", "squaredcos_cap_v2"),
                       "num_train_timesteps": (1000,), "solver_order": (2,), "solver_type": ("midpoint",),
                       "timestep_spacing": ("linspace",), "final_sigmas_type": ("zero",), "use_karras_sigmas": (False,),
                       "lower_order_final": (True,), "thresholding": (False,),
                       "algorithm_type": ("dpmsolver++", "sde-dpmsolver++")}


def check_scheduler(scheduler) -> None:
    """Raise VVError unless the model's noise scheduler is configured the way this library's solver is.  Fields the
    config object does not carry are skipped (the reference's defaults are the implemented ones)."""
    cfg = getattr(scheduler, "config", None)
    if cfg is None:
        return
    get = cfg.get if isinstance(cfg, dict) else (lambda k, d=None: getattr(cfg, k, d))
    for key, allowed in _SCHEDULER_EXPECTED.items():
        val = get(key, None)
        if val is not None and val not in allowed:
            raise _lib.VVError(f"vibevoice_b200: noise_scheduler.config.{key} = {val!r} is not implemented "
                               f"(supported: {', '.join(map(repr, allowed))}); refusing to produce wrong latents")


def install(model, max_batch: int = 64, max_offline_rows: Optional[int] = None, frame_tail: bool = False,
            prefill: bool = False, prefill_rows: int = 512, fixed_shape: bool = True) -> Installed:
    """Rebind the acoustic path of ``model`` to libvvb200.  max_batch: largest number of batch rows
    generate() will carry (cache slots; the head is sized for 2*max_batch CFG rows).  frame_tail: also rebind
    the semantic-tokenizer encode and the two SpeechConnectors (the rest of the frame iteration).  prefill: also
    rebind the voice-prompt path (acoustic_tokenizer.encode without a cache and _process_speech_inputs);
    prefill_rows: frames per internal time chunk of the prompt encoder (scratch is about 2 MB per frame).
    fixed_shape: run the head at max_batch rows whatever the number of active rows (no mid-generation program builds)."""
    _lib.load()   # raises if the CUDA library is missing: no fallback
    core = model.model
    p = next(core.prediction_head.parameters())
    if p.device.type != "cuda":
        raise _lib.VVError("vibevoice_b200.install: the model must live on a CUDA device (no CPU fallback)")
    check_scheduler(getattr(core, "noise_scheduler", None))
    head = DiffusionHeadB200.from_module(core.prediction_head, max_rows=2 * max_batch)
    if fixed_shape:
        # generate() hands sample_speech_tokens only the rows in diffusion mode, a count that changes whenever a row
        # finishes (modeling_vibevoice_inference.py:588-590): run the head at ONE shape, max_batch rows, so that the
        # library never builds a program (and never host-synchronises) in the middle of a generation
        head.fixed_rows = int(max_batch)
    dec = AcousticDecoderB200.from_module(core.acoustic_tokenizer, max_rows=max_offline_rows or max(max_batch, 64))
    inst = Installed(model, head, dec, max_batch)
    latent = int(model.config.acoustic_vae_dim)

    @torch.no_grad()
    def sample_speech_tokens(self, condition, neg_condition, cfg_scale=3.0):
        # noise: the reference's exact draw, CPU global generator (modeling_vibevoice_inference.py:716)
        n = condition.shape[0]
        if n > max_batch:
            raise _lib.VVError(f"vibevoice_b200: {n} rows in diffusion mode exceed install(max_batch={max_batch})")
        check_scheduler(self.model.noise_scheduler)
        noise = torch.randn(2 * n, latent)
        return head.sample_speech_tokens(condition, neg_condition, cfg_scale,
                                         num_steps=int(self.ddpm_inference_steps),
                                         algorithm_type=_algorithm_type(self.model.noise_scheduler), noise=noise)

    def head_forward(noisy_images, timesteps, condition):
        return head.forward(noisy_images, timesteps, condition)

    @torch.no_grad()
    def decode(latents, cache=None, sample_indices=None, use_cache=False, debug=False):
        if use_cache and cache is not None:
            return dec.decode(latents, cache=inst.arena_for(cache), sample_indices=sample_indices, use_cache=True)
        return dec.decode(latents)

    inst._bind(model, "sample_speech_tokens", types.MethodType(sample_speech_tokens, model))
    inst._bind(core.prediction_head, "forward", head_forward)
    inst._bind(core.acoustic_tokenizer, "decode", decode)
    if frame_tail:
        enc = SemanticEncoderB200.from_module(core.semantic_tokenizer, max_rows=max(max_batch, 64))
        conn_a = SpeechConnectorB200.from_module(core.acoustic_connector, max_rows=max(max_batch, 64))
        conn_s = SpeechConnectorB200.from_module(core.semantic_connector, max_rows=max(max_batch, 64))
        inst.enc, inst.conn_a, inst.conn_s = enc, conn_a, conn_s

        @torch.no_grad()
        def encode(audio, cache=None, sample_indices=None, use_cache=False, debug=False):
            if use_cache and cache is not None:
                return enc.encode(audio, cache=inst.arena_for(cache, enc), sample_indices=sample_indices, use_cache=True)
            return enc.encode(audio)

        def acoustic_connector(features, **kwargs):
            return conn_a(features)

        def semantic_connector(features, **kwargs):
            return conn_s(features)

        inst._bind(core.semantic_tokenizer, "encode", encode)
        inst._bind(core.acoustic_connector, "forward", acoustic_connector)
        inst._bind(core.semantic_connector, "forward", semantic_connector)
    if prefill:
        rows = max(int(prefill_rows), max_batch)
        aenc = AcousticEncoderB200.from_module(core.acoustic_tokenizer, max_rows=rows)
        conn_p = SpeechConnectorB200.from_module(core.acoustic_connector, max_rows=max(rows, 64))
        inst.aenc, inst.conn_p = aenc, conn_p
        ref_encode = core.acoustic_tokenizer.encode          # streaming encode of the acoustic tokenizer: torch path

        @torch.no_grad()
        def acoustic_encode(audio, cache=None, sample_indices=None, use_cache=False, debug=False):
            if use_cache and cache is not None:
                return ref_encode(audio, cache=cache, sample_indices=sample_indices, use_cache=True, debug=debug)
            return aenc.encode(audio)

        @torch.no_grad()
        def _process_speech_inputs(self, speech_tensors, speech_masks, speech_type="audio"):
            if speech_type == "audio":
                out = aenc.encode(speech_tensors.unsqueeze(1))
            elif speech_type == "pt":
                out = EncoderOutputB200(speech_tensors, aenc.fix_std)
            else:
                raise NotImplementedError(f"Speech type {speech_type} not implemented")
            feats = out.sample_features(aenc.std_dist_type, self.model.speech_bias_factor,
                                        self.model.speech_scaling_factor)
            return feats, conn_p(feats)[speech_masks.cpu()]

        inst._bind(core.acoustic_tokenizer, "encode", acoustic_encode)
        inst._bind(model, "_process_speech_inputs", types.MethodType(_process_speech_inputs, model))
    return inst
