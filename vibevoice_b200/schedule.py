"""Host-side DPM-Solver++ schedule tables (product code; never imports oracle/).

Mirrors DPMSolverMultistepScheduler as VibeVoice instantiates it
(modeling_vibevoice.py:138-142: num_train_timesteps=1000, beta_schedule="cosine",
prediction_type="v_prediction"; defaults solver_order=2, midpoint, lower_order_final,
final_sigmas_type="zero", timestep_spacing="linspace" — schedule/dpm_solver.py:203-227).

The reference recomputes these tables with numpy on the host for EVERY frame
(set_timesteps, schedule/dpm_solver.py:321-423); here they are computed once per
(num_steps, dtype) and uploaded with vv_head_set_schedule.
"""
from __future__ import annotations

import functools
import math

import numpy as np
import torch

ALGOS = {"dpmsolver++": 0, "sde-dpmsolver++": 1}


@functools.lru_cache(maxsize=None)
def _alphas_cumprod(num_train_timesteps: int = 1000) -> np.ndarray:
    # betas_for_alpha_bar, cosine (schedule/dpm_solver.py:28-83): betas in python doubles,
    # stored fp32; cumprod in fp32 (:252-253).
    def alpha_bar(t):
        return math.cos((t + 0.008) / 1.008 * math.pi / 2) ** 2

    betas = [min(1 - alpha_bar((i + 1) / num_train_timesteps) / alpha_bar(i / num_train_timesteps), 0.999)
             for i in range(num_train_timesteps)]
    betas = torch.tensor(betas, dtype=torch.float32)
    return torch.cumprod(1.0 - betas, dim=0).numpy()


@functools.lru_cache(maxsize=None)
def timesteps_and_sigmas(num_inference_steps: int, num_train_timesteps: int = 1000):
    """(timesteps int64 [S], sigmas float32 [S+1]) — set_timesteps, 'linspace', final sigma 0
    (schedule/dpm_solver.py:358-364, 384, 397, 401-408)."""
    if num_inference_steps < 1:
        raise ValueError("num_inference_steps must be >= 1")
    ac = torch.from_numpy(_alphas_cumprod(num_train_timesteps))
    ts = np.linspace(0, num_train_timesteps - 1, num_inference_steps + 1).round()[::-1][:-1].copy().astype(np.int64)
    all_sigmas = (((1 - ac) / ac) ** 0.5).numpy()
    sig = np.interp(ts, np.arange(0, len(all_sigmas)), all_sigmas)
    sig = np.concatenate([sig, [0]]).astype(np.float32)
    return ts, sig


def timestep_values(timesteps: np.ndarray, model_dtype: torch.dtype) -> np.ndarray:
    """The value the head actually sees: `t.repeat(..).to(combined)` casts the int64 timestep to
    the model dtype (modeling_vibevoice_inference.py:720) — bf16 maps 999 -> 1000, 899 -> 900 ..."""
    return torch.from_numpy(np.asarray(timesteps)).to(model_dtype).to(torch.float32).numpy().copy()
