"""TEST INFRASTRUCTURE — not part of the product path.

Imports the *unmodified* reference modules from /root/reference in this container so the
oracle restatement (oracle/vv_oracle.py) can be pinned against them and golden vectors can
be generated (tests/golden/make_golden.py).  /root/reference does not exist on the GPU box:
nothing under ``-m gpu``, ``smoke()`` or ``bench.py`` may call this module.

Two shims are needed:
  1. transformers 5.5 already ships ``vibevoice_acoustic_tokenizer``; the reference's
     ``AutoModel.register(...)`` at import raises, so registration is made a no-op
     (modular_vibevoice_tokenizer.py:1188-1189, modular_vibevoice_diffusion_head.py:283).
  2. ``diffusers`` is not installed; vibevoice/schedule/dpm_solver.py:23-26 imports seven
     names from it, none of which does arithmetic.  A stub provides them.
"""
from __future__ import annotations

import enum
import functools
import inspect
import os
import sys
import types
from dataclasses import dataclass

REFERENCE_ROOT = os.environ.get("VV_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "vibevoice", "modular"))


def _install_diffusers_stub() -> None:
    if "diffusers" in sys.modules:
        return
    import torch

    class _Config(types.SimpleNamespace):
        def __getitem__(self, k):
            return getattr(self, k)

        def get(self, k, d=None):
            return getattr(self, k, d)

    class ConfigMixin:
        def register_to_config(self, **kw):
            if not hasattr(self, "config"):
                self.config = _Config()
            for k, v in kw.items():
                setattr(self.config, k, v)

        @classmethod
        def from_config(cls, config, **overrides):
            kw = dict(vars(config))
            kw.update(overrides)
            sig = inspect.signature(cls.__init__).parameters
            return cls(**{k: v for k, v in kw.items() if k in sig})

    def register_to_config(init):
        @functools.wraps(init)
        def wrapper(self, *args, **kwargs):
            sig = inspect.signature(init)
            bound = sig.bind(self, *args, **kwargs)
            bound.apply_defaults()
            cfg = {k: v for k, v in bound.arguments.items() if k != "self"}
            self.config = _Config(**cfg)
            init(self, *args, **kwargs)

        return wrapper

    def deprecate(*a, **k):
        return None

    def randn_tensor(shape, generator=None, device=None, dtype=None, layout=None):
        return torch.randn(shape, generator=generator, device=device, dtype=dtype)

    class KarrasDiffusionSchedulers(enum.Enum):
        DPMSolverMultistepScheduler = 1

    class SchedulerMixin:
        pass

    @dataclass
    class SchedulerOutput:
        prev_sample: "torch.Tensor"

    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    mod("diffusers")
    mod("diffusers.configuration_utils", ConfigMixin=ConfigMixin, register_to_config=register_to_config)
    mod("diffusers.utils", deprecate=deprecate)
    mod("diffusers.utils.torch_utils", randn_tensor=randn_tensor)
    mod("diffusers.schedulers")
    mod("diffusers.schedulers.scheduling_utils", KarrasDiffusionSchedulers=KarrasDiffusionSchedulers,
        SchedulerMixin=SchedulerMixin, SchedulerOutput=SchedulerOutput)


_loaded = None


def load():
    """Returns a namespace with the reference classes on the hot path."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    sys.dont_write_bytecode = True
    from transformers.models.auto import AutoModel, AutoModelForCausalLM

    AutoModel.register = classmethod(lambda cls, *a, **k: None)
    AutoModelForCausalLM.register = classmethod(lambda cls, *a, **k: None)
    _install_diffusers_stub()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    from vibevoice.modular.configuration_vibevoice import (
        VibeVoiceAcousticTokenizerConfig,
        VibeVoiceDiffusionHeadConfig,
        VibeVoiceSemanticTokenizerConfig,
    )
    from vibevoice.modular.modular_vibevoice_diffusion_head import VibeVoiceDiffusionHead
    from vibevoice.modular.modular_vibevoice_tokenizer import (
        VibeVoiceAcousticTokenizerModel,
        VibeVoiceSemanticTokenizerModel,
        VibeVoiceTokenizerStreamingCache,
    )
    from vibevoice.schedule.dpm_solver import DPMSolverMultistepScheduler

    _loaded = types.SimpleNamespace(
        HeadConfig=VibeVoiceDiffusionHeadConfig,
        AcousticConfig=VibeVoiceAcousticTokenizerConfig,
        Head=VibeVoiceDiffusionHead,
        Acoustic=VibeVoiceAcousticTokenizerModel,
        Cache=VibeVoiceTokenizerStreamingCache,
        Scheduler=DPMSolverMultistepScheduler,
        SemanticConfig=VibeVoiceSemanticTokenizerConfig,
        Semantic=VibeVoiceSemanticTokenizerModel,
    )
    return _loaded


def make_semantic(cfg, W):
    """Reference VibeVoiceSemanticTokenizerModel with the weights W (TokenizerEncoder state-dict names).
    cfg.ratios is in encoder order; the config wants them reversed (modular_vibevoice_tokenizer.py:701)."""
    ref = load()
    sc = ref.SemanticConfig(vae_dim=cfg.dim, encoder_n_filters=cfg.n_filters,
                            encoder_ratios=list(reversed(cfg.ratios)),
                            encoder_depths="-".join(str(d) for d in cfg.depths), layernorm_eps=cfg.eps)
    model = ref.Semantic(sc).eval()
    res = model.load_state_dict({"encoder." + k: v for k, v in W.items()}, strict=True)
    assert not res.unexpected_keys and not res.missing_keys
    return model


def make_acoustic_encoder(cfg, W, fix_std=0.5, std_dist_type="gaussian"):
    """Reference VibeVoiceAcousticTokenizerModel whose ENCODER carries the weights W (TokenizerEncoder state-dict
    names); the decoder keeps its random initialisation (the prefill path, modeling_vibevoice_inference.py:150-164,
    does not touch it)."""
    ref = load()
    ac = ref.AcousticConfig(vae_dim=cfg.dim, encoder_n_filters=cfg.n_filters, decoder_n_filters=cfg.n_filters,
                            encoder_ratios=list(reversed(cfg.ratios)),
                            encoder_depths="-".join(str(d) for d in cfg.depths), layernorm_eps=cfg.eps,
                            fix_std=fix_std, std_dist_type=std_dist_type)
    model = ref.Acoustic(ac).eval()
    res = model.load_state_dict({"encoder." + k: v for k, v in W.items()}, strict=False)
    assert not res.unexpected_keys
    assert all(k.startswith("decoder.") for k in res.missing_keys)
    return model


def encoder_output(mean, std):
    """The reference's VibeVoiceTokenizerEncoderOutput (modular_vibevoice_tokenizer.py:954-1000)."""
    load()
    from vibevoice.modular.modular_vibevoice_tokenizer import VibeVoiceTokenizerEncoderOutput

    return VibeVoiceTokenizerEncoderOutput(mean=mean, std=std)


def make_connector(cfg, W):
    """Reference SpeechConnector (modeling_vibevoice.py:58-69).  Importing modeling_vibevoice pulls the whole
    model file in; it only needs the shims load() already installed."""
    load()
    from vibevoice.modular.modeling_vibevoice import SpeechConnector

    m = SpeechConnector(cfg.input_dim, cfg.output_dim).eval()
    m.load_state_dict(W, strict=True)
    return m


def make_scheduler(algorithm_type: str = "dpmsolver++"):
    """Scheduler exactly as VibeVoiceModel builds it (modeling_vibevoice.py:138-142)."""
    ref = load()
    return ref.Scheduler(num_train_timesteps=1000, beta_schedule="cosine",
                         prediction_type="v_prediction", algorithm_type=algorithm_type)


def ref_sample_speech_tokens(head, scheduler, cond, neg, cfg_scale, steps, noise=None):
    """modeling_vibevoice_inference.py:712-725 with self.model.* replaced by the two objects.

    ``noise`` (optional [2n, L]) replaces the torch.randn draw so the oracle/CUDA path can be
    fed the identical tensor; when None the reference's own draw is used.
    """
    import torch

    with torch.no_grad():
        scheduler.set_timesteps(steps)
        condition = torch.cat([cond, neg], dim=0)
        if noise is None:
            speech = torch.randn(condition.shape[0], head.config.latent_size).to(condition)
        else:
            speech = noise.to(condition)
        for t in scheduler.timesteps:
            half = speech[: len(speech) // 2]
            combined = torch.cat([half, half], dim=0)
            eps = head(combined, t.repeat(combined.shape[0]).to(combined), condition=condition)
            cond_eps, uncond_eps = torch.split(eps, len(eps) // 2, dim=0)
            half_eps = uncond_eps + cfg_scale * (cond_eps - uncond_eps)
            eps = torch.cat([half_eps, half_eps], dim=0)
            speech = scheduler.step(eps, t, speech).prev_sample
        return speech[: len(speech) // 2]
