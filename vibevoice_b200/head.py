"""Host mirror of VibeVoiceDiffusionHead + sample_speech_tokens over libvvb200.

Same names / argument meaning as the reference:
  * ``DiffusionHeadB200.forward(noisy_images, timesteps, condition)``
        <- VibeVoiceDiffusionHead.forward (modular_vibevoice_diffusion_head.py:254-259)
  * ``DiffusionHeadB200.sample_speech_tokens(condition, neg_condition, cfg_scale, ...)``
        <- VibeVoiceForConditionalGenerationInference.sample_speech_tokens
           (modeling_vibevoice_inference.py:712-725)
PyTorch only provides device memory and the stream; all arithmetic runs in the library.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import numpy as np
import torch

from . import _lib, schedule, synth
from ._lib import check, current_stream, ptr


class DiffusionHeadB200:
    def __init__(self, weights: Dict[str, torch.Tensor], cfg: synth.HeadCfg, device="cuda",
                 max_rows: int = 256, model_dtype: torch.dtype = torch.bfloat16):
        """weights: reference state-dict (name -> tensor, any float dtype, any device)."""
        self.lib = _lib.load()
        self.cfg = cfg
        self.device = torch.device(device)
        self.model_dtype = model_dtype
        self.max_rows = int(max_rows)
        staged = [w.detach().to(self.device, torch.float32).contiguous()
                  for w in synth.ordered_weights(synth.head_param_shapes(cfg), weights, what="head weight")]
        arr = (C.c_void_p * len(staged))(*[t.data_ptr() for t in staged])
        desc = _lib.VVHeadDesc(cfg.hidden, cfg.ffn, cfg.latent, cfg.layers, cfg.freq_dim, cfg.eps,
                               _lib.VV_DTYPE_F32, self.max_rows)
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            torch.cuda.synchronize()
            check(self.lib.vv_head_create(C.byref(desc), arr, len(staged), C.byref(h)), "vv_head_create")
        self._h = h
        del staged
        self._sched_key = None
        # fixed launch shape (dropin.install): every sample_speech_tokens call runs `fixed_rows` rows, the spare ones on
        # zeros - one persistent program for the whole generation instead of one (and a host sync) per active-row count
        self.fixed_rows: Optional[int] = None
        self._pad = None

    @staticmethod
    def spec_from_module(module):
        """(state dict, HeadCfg) of a live reference ``VibeVoiceDiffusionHead`` - host only, no CUDA."""
        c = module.config
        cfg = synth.HeadCfg(hidden=c.hidden_size, ffn_ratio=c.head_ffn_ratio, latent=c.latent_size,
                            layers=c.head_layers, eps=c.rms_norm_eps)
        return dict(module.state_dict()), cfg

    @classmethod
    def from_module(cls, module, **kw):
        """Build from a live reference ``VibeVoiceDiffusionHead`` (reads config + state_dict)."""
        W, cfg = cls.spec_from_module(module)
        p = next(module.parameters())
        kw.setdefault("device", p.device)
        kw.setdefault("model_dtype", p.dtype)
        return cls(W, cfg, **kw)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            self.lib.vv_head_destroy(h)
            self._h = None

    # ---- schedule -----------------------------------------------------------------------
    def set_schedule(self, num_steps: int, algorithm_type: str = "dpmsolver++") -> None:
        """Once per (steps, algorithm, dtype) instead of noise_scheduler.set_timesteps per frame."""
        key = (int(num_steps), algorithm_type, self.model_dtype)
        if key == self._sched_key:
            return
        if algorithm_type not in schedule.ALGOS:
            raise ValueError(f"unsupported algorithm_type {algorithm_type!r}")
        ts, sig = schedule.timesteps_and_sigmas(int(num_steps))
        tv = np.ascontiguousarray(schedule.timestep_values(ts, self.model_dtype), dtype=np.float32)
        sig = np.ascontiguousarray(sig, dtype=np.float32)
        with torch.cuda.device(self.device):
            check(self.lib.vv_head_set_schedule(
                self._h, len(tv), tv.ctypes.data_as(C.POINTER(C.c_float)),
                sig.ctypes.data_as(C.POINTER(C.c_float)), schedule.ALGOS[algorithm_type],
                current_stream()), "vv_head_set_schedule")
        self._sched_key = key
        self.num_steps = int(num_steps)
        self.algorithm_type = algorithm_type

    # ---- drop-ins -----------------------------------------------------------------------
    def _io(self, t: torch.Tensor):
        if t.dtype == torch.bfloat16:
            return t.contiguous(), _lib.VV_DTYPE_BF16
        return t.to(torch.float32).contiguous(), _lib.VV_DTYPE_F32

    @torch.no_grad()
    def forward(self, noisy_images: torch.Tensor, timesteps: torch.Tensor, condition: torch.Tensor):
        M = noisy_images.shape[0]
        z, io = self._io(noisy_images.to(self.device))
        cond = condition.to(self.device, z.dtype).contiguous()
        t = timesteps.to(self.device).to(torch.float32).contiguous()
        out = torch.empty(M, self.cfg.latent, device=self.device, dtype=z.dtype)
        with torch.cuda.device(self.device):
            check(self.lib.vv_head_forward(self._h, ptr(z), ptr(t), ptr(cond), ptr(out), M, io,
                                           current_stream()), "vv_head_forward")
        return out.to(noisy_images.dtype)

    __call__ = forward

    @torch.no_grad()
    def sample_speech_tokens(self, condition: torch.Tensor, neg_condition: torch.Tensor,
                             cfg_scale: float = 3.0, num_steps: Optional[int] = None,
                             algorithm_type: Optional[str] = None,
                             noise: Optional[torch.Tensor] = None,
                             sde_noise: Optional[torch.Tensor] = None) -> torch.Tensor:
        """[n,H],[n,H] -> [n,L].  noise defaults to the reference's own draw:
        ``torch.randn(2n, L)`` from the CPU global generator (modeling_vibevoice_inference.py:716)."""
        if num_steps is not None or algorithm_type is not None or self._sched_key is None:
            self.set_schedule(num_steps or getattr(self, "num_steps", 10),
                              algorithm_type or getattr(self, "algorithm_type", "dpmsolver++"))
        n = condition.shape[0]
        L = self.cfg.latent
        cond, io = self._io(condition.to(self.device))
        neg = neg_condition.to(self.device, cond.dtype).contiguous()
        if noise is None:
            noise = torch.randn(2 * n, L)
        noise = noise.to(self.device, torch.float32, non_blocking=True).contiguous()
        if self.algorithm_type == "sde-dpmsolver++":
            if sde_noise is None:  # randn_tensor(..., device=model_output.device) per step (dpm_solver.py:994-997)
                sde_noise = torch.randn(self.num_steps, n, L, device=self.device, dtype=torch.float32)
            sde_noise = sde_noise.to(self.device, torch.float32).contiguous()
            assert sde_noise.shape == (self.num_steps, n, L)
        else:
            sde_noise = None
        n_run = n
        if self.fixed_rows is not None and n < self.fixed_rows:
            # pad to the fixed shape: rows are independent everywhere in the head, the spare rows' outputs are dropped.
            # The noise of the active rows stays the caller's draw (rows 0..n-1 of torch.randn(2n, L), :716-718).
            n_run = int(self.fixed_rows)
            key = (n_run, cond.dtype)
            if self._pad is None or self._pad[0] != key:
                z = lambda *shape, dt=cond.dtype: torch.zeros(*shape, device=self.device, dtype=dt)   # noqa: E731
                self._pad = (key, z(n_run, cond.shape[1]), z(n_run, cond.shape[1]), z(2 * n_run, L, dt=torch.float32),
                             z(self.num_steps, n_run, L, dt=torch.float32))
            _, pc, pn, pz, ps = self._pad
            pc[:n].copy_(cond); pn[:n].copy_(neg); pz[:n].copy_(noise[:n])
            if sde_noise is not None:
                if ps.shape[0] != self.num_steps:
                    ps = torch.zeros(self.num_steps, n_run, L, device=self.device, dtype=torch.float32)
                    self._pad = (key, pc, pn, pz, ps)
                ps[:, :n].copy_(sde_noise)
                sde_noise = ps
            cond, neg, noise = pc, pn, pz
        out = torch.empty(n_run, L, device=self.device, dtype=cond.dtype)
        with torch.cuda.device(self.device):
            check(self.lib.vv_head_sample(self._h, ptr(cond), ptr(neg), ptr(noise), ptr(sde_noise),
                                          float(cfg_scale), ptr(out), n_run, io, current_stream()),
                  "vv_head_sample")
        return out[:n].to(condition.dtype)
