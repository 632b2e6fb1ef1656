"""TEST INFRASTRUCTURE — CPU restatement (torch fp32/fp64 on the host) of the reference's
per-frame acoustic generation path.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this; the product (vibevoice_b200/) never
does.

Parity pinning: the reference ships no tests or golden vectors for this path,
so this oracle is pinned against the *live reference modules* imported from /root/reference
(tests/test_oracle_vs_reference.py, CPU container only) and against fixtures those modules
generated (tests/golden/, made by tests/golden/make_golden.py).  On the GPU box the reference
tree does not exist; the oracle + fixtures stand in for it.

Every function cites the reference lines it restates (paths relative to the reference repo).
``emulate_bf16=True`` additionally rounds to bf16 at exactly the points where the CUDA
pipeline stores bf16 (weights, GEMM operands, decoder residual stream), so the CUDA path can
be compared against it with a tight tolerance; with False it is the fp32 "truth".
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.nn.functional as F

# --------------------------------------------------------------------------------------
# Schedule: DPMSolverMultistepScheduler(num_train_timesteps=1000, beta_schedule="cosine",
#   prediction_type="v_prediction")  (modeling_vibevoice.py:138-142)
# --------------------------------------------------------------------------------------


def cosine_alphas_cumprod(num_train_timesteps: int = 1000, max_beta: float = 0.999) -> torch.Tensor:
    """betas_for_alpha_bar (schedule/dpm_solver.py:28-83) + cumprod (:252-253), fp32."""

    def alpha_bar(t):
        return math.cos((t + 0.008) / 1.008 * math.pi / 2) ** 2

    betas = []
    for i in range(num_train_timesteps):
        t1 = i / num_train_timesteps
        t2 = (i + 1) / num_train_timesteps
        betas.append(min(1 - alpha_bar(t2) / alpha_bar(t1), max_beta))
    betas = torch.tensor(betas, dtype=torch.float32)
    return torch.cumprod(1.0 - betas, dim=0)


def set_timesteps(num_inference_steps: int, num_train_timesteps: int = 1000):
    """set_timesteps, 'linspace' spacing, final_sigmas_type='zero' (schedule/dpm_solver.py:358-408).

    Returns (timesteps int64 [S], sigmas float32 [S+1]).  lambda_min_clipped = -inf, so
    last_timestep = num_train_timesteps (:354-355).
    """
    ac = cosine_alphas_cumprod(num_train_timesteps)
    last_timestep = num_train_timesteps
    timesteps = (
        np.linspace(0, last_timestep - 1, num_inference_steps + 1).round()[::-1][:-1].copy().astype(np.int64)
    )
    sigmas = (((1 - ac) / ac) ** 0.5).numpy()
    sigmas = np.interp(timesteps, np.arange(0, len(sigmas)), sigmas)
    sigmas = np.concatenate([sigmas, [0]]).astype(np.float32)
    return timesteps, sigmas


def round_timesteps(timesteps: np.ndarray, dtype: torch.dtype) -> np.ndarray:
    """`t.repeat(..).to(combined)` casts the int64 timestep to the model dtype
    (modeling_vibevoice_inference.py:720): bf16 turns 999 into 1000, 899 into 900, ..."""
    return torch.from_numpy(timesteps).to(dtype).to(torch.float32).numpy()


def _alpha_sigma(sigma: torch.Tensor):
    """_sigma_to_alpha_sigma_t (schedule/dpm_solver.py:483-487)."""
    alpha_t = 1 / ((sigma ** 2 + 1) ** 0.5)
    return alpha_t, sigma * alpha_t


def solver_coefficients(sigmas: np.ndarray, algo: str = "dpmsolver++") -> List[Dict[str, float]]:
    """Per-step scalars of step()/first_order/second_order (schedule/dpm_solver.py:935-1022,
    669-686, 738-793), evaluated with fp32 0-dim tensors exactly as the reference does.

    Step i: x0 = alpha_s * x - sigma_s * v                               (:581-584)
      order 1 (i == 0 or i == S-1):  x' = cx * x + c0 * x0 (+ cn * noise)
      order 2:  D1 = (x0 - x0_prev) * inv_r0 ;  x' = cx * x + c0 * x0 + c1 * D1 (+ cn * noise)
    """
    s = torch.from_numpy(np.asarray(sigmas, dtype=np.float32))
    S = len(s) - 1
    out = []
    for i in range(S):
        first = i == 0
        # lower_order_final: final_sigmas_type == "zero" makes the last step first-order (:978-982)
        last = i == S - 1
        order = 1 if (first or last) else 2
        alpha_s0, sigma_s0 = _alpha_sigma(s[i])
        alpha_t, sigma_t = _alpha_sigma(s[i + 1])
        lambda_t = torch.log(alpha_t) - torch.log(sigma_t)
        lambda_s0 = torch.log(alpha_s0) - torch.log(sigma_s0)
        h = lambda_t - lambda_s0
        c = {"alpha_s": float(alpha_s0), "sigma_s": float(sigma_s0), "order": order,
             "c1": 0.0, "inv_r0": 0.0, "cn": 0.0}
        if order == 2:
            alpha_s1, sigma_s1 = _alpha_sigma(s[i - 1])
            lambda_s1 = torch.log(alpha_s1) - torch.log(sigma_s1)
            h_0 = lambda_s0 - lambda_s1
            r0 = h_0 / h
            c["inv_r0"] = float(1.0 / r0)
        if algo == "dpmsolver++":
            c["cx"] = float(sigma_t / sigma_s0)
            c["c0"] = float(-(alpha_t * (torch.exp(-h) - 1.0)))
            if order == 2:
                c["c1"] = float(-0.5 * (alpha_t * (torch.exp(-h) - 1.0)))
        elif algo == "sde-dpmsolver++":
            c["cx"] = float(sigma_t / sigma_s0 * torch.exp(-h))
            c["c0"] = float(alpha_t * (1 - torch.exp(-2.0 * h)))
            if order == 2:
                c["c1"] = float(0.5 * (alpha_t * (1 - torch.exp(-2.0 * h))))
            c["cn"] = float(sigma_t * torch.sqrt(1.0 - torch.exp(-2 * h)))
        else:
            raise ValueError(algo)
        out.append(c)
    return out


# --------------------------------------------------------------------------------------
# Diffusion head  (modular/modular_vibevoice_diffusion_head.py)
# --------------------------------------------------------------------------------------


def _bf(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.bfloat16).to(t.dtype)


def _ident(t: torch.Tensor) -> torch.Tensor:
    return t


def rmsnorm(x: torch.Tensor, weight: Optional[torch.Tensor], eps: float) -> torch.Tensor:
    """RMSNorm.forward (modular_vibevoice_diffusion_head.py:31-38): fp32 normalise, then * weight."""
    y = x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + eps)
    return y * weight if weight is not None else y


def timestep_embedding(t: torch.Tensor, dim: int, max_period: float = 10000.0) -> torch.Tensor:
    """TimestepEmbedder.timestep_embedding (modular_vibevoice_diffusion_head.py:66-88): cat[cos, sin]."""
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(0, half, dtype=torch.float32) / half)
    args = t[:, None].float() * freqs[None]
    return torch.cat([torch.cos(args), torch.sin(args)], dim=-1).to(t.dtype)


def t_embed(W: Dict[str, torch.Tensor], t: torch.Tensor, freq_dim: int, r=_ident) -> torch.Tensor:
    """TimestepEmbedder.forward (:90-93): Linear -> SiLU -> Linear, no biases."""
    e = r(timestep_embedding(t, freq_dim))
    h = r(F.silu(F.linear(e, r(W["t_embedder.mlp.0.weight"]))))
    return F.linear(h, r(W["t_embedder.mlp.2.weight"]))


def head_forward(W: Dict[str, torch.Tensor], z: torch.Tensor, t: torch.Tensor, cond: torch.Tensor,
                 layers: int = 4, eps: float = 1e-5, freq_dim: int = 256,
                 emulate_bf16: bool = False) -> torch.Tensor:
    """VibeVoiceDiffusionHead.forward (modular_vibevoice_diffusion_head.py:254-280).

    z [M,L], t [M] (float, already rounded to the model dtype by the caller), cond [M,H].
    With emulate_bf16 the GEMM operands (weights and activations entering a Linear) and the
    adaLN outputs are rounded to bf16 as in the CUDA pipeline; the residual stream stays fp32.
    """
    r = _bf if emulate_bf16 else _ident
    x = F.linear(r(z), r(W["noisy_images_proj.weight"]))                       # :271
    temb = t_embed(W, t, freq_dim, r)                                          # :272
    c = F.linear(r(cond), r(W["cond_proj.weight"])) + temb                     # :273-274
    sc = r(F.silu(c))                                                          # adaLN_modulation[0]
    for i in range(layers):                                                    # HeadLayer.forward :158-161
        p = f"layers.{i}."
        mod = r(F.linear(sc, r(W[p + "adaLN_modulation.1.weight"])))
        shift, scale, gate = mod.chunk(3, dim=-1)
        y = r(rmsnorm(x, W[p + "norm.weight"], eps) * (1 + scale) + shift)     # modulate :43-45
        g = F.linear(y, r(W[p + "ffn.gate_proj.weight"]))
        u = F.linear(y, r(W[p + "ffn.up_proj.weight"]))
        hmid = r(F.silu(g) * u)                                                # FeedForwardNetwork :116-123
        x = x + gate * F.linear(hmid, r(W[p + "ffn.down_proj.weight"]))
    mod = r(F.linear(sc, r(W["final_layer.adaLN_modulation.1.weight"])))       # FinalLayer.forward :184-188
    shift, scale = mod.chunk(2, dim=-1)
    y = r(rmsnorm(x, None, eps) * (1 + scale) + shift)
    return F.linear(y, r(W["final_layer.linear.weight"]))


def sample_speech_tokens(W: Dict[str, torch.Tensor], cond: torch.Tensor, neg: torch.Tensor,
                         cfg_scale: float, noise: torch.Tensor, steps: int = 10,
                         algo: str = "dpmsolver++", sde_noise: Optional[torch.Tensor] = None,
                         t_dtype: torch.dtype = torch.bfloat16, layers: int = 4, eps: float = 1e-5,
                         freq_dim: int = 256, emulate_bf16: bool = False,
                         return_trajectory: bool = False):
    """sample_speech_tokens (modeling_vibevoice_inference.py:712-725) + scheduler.step
    (schedule/dpm_solver.py:935-1022).

    noise: [2n, L], the torch.randn draw of :716 (rows n.. are carried but never read).
    sde_noise: [S, 2n, L] for 'sde-dpmsolver++' (the per-step randn_tensor of :994-997).
    t_dtype: dtype the timesteps are cast to before the sinusoid (:720); bf16 on GPU, fp32 on CPU.
    Only rows :n of the state are tracked (the reference updates rows n: too and discards them).
    """
    n = cond.shape[0]
    timesteps, sigmas = set_timesteps(steps)
    tvals = torch.from_numpy(round_timesteps(timesteps, t_dtype))
    coefs = solver_coefficients(sigmas, algo)
    condition = torch.cat([cond, neg], dim=0)                                  # :715
    x = noise[:n].clone().float()                                              # speech[:n]  :716-718
    prev_x0 = None
    traj = []
    for i in range(steps):
        combined = torch.cat([x, x], dim=0)                                    # :719
        tt = tvals[i].repeat(2 * n)
        v = head_forward(W, combined, tt, condition, layers, eps, freq_dim, emulate_bf16)  # :720
        c_eps, u_eps = v[:n], v[n:]                                            # :721
        v = u_eps + cfg_scale * (c_eps - u_eps)                                # :722
        k = coefs[i]
        x0 = k["alpha_s"] * x - k["sigma_s"] * v                               # convert_model_output :581-584
        nz = sde_noise[i][:n].float() if (algo == "sde-dpmsolver++") else 0.0
        if k["order"] == 1:                                                    # :1003-1004
            xn = k["cx"] * x + k["c0"] * x0 + (k["cn"] * nz if k["cn"] != 0.0 else 0.0)
        else:                                                                  # :1005-1006, :752-764
            d1 = k["inv_r0"] * (x0 - prev_x0)
            xn = k["cx"] * x + k["c0"] * x0 + k["c1"] * d1 + (k["cn"] * nz if k["cn"] != 0.0 else 0.0)
        prev_x0 = x0
        x = xn
        if return_trajectory:
            traj.append(x.clone())
    return (x, traj) if return_trajectory else x


# --------------------------------------------------------------------------------------
# Acoustic tokenizer decoder + streaming cache  (modular/modular_vibevoice_tokenizer.py)
# --------------------------------------------------------------------------------------


def conv_rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    """ConvRMSNorm.forward (:77-91): RMS over channels per time step; x is [n, C, T]."""
    xt = x.transpose(1, 2)
    y = xt * torch.rsqrt(xt.pow(2).mean(-1, keepdim=True) + eps)
    return (y * weight).transpose(1, 2)


class DecoderState:
    """Functional stand-in for VibeVoiceTokenizerStreamingCache (:193-256) with per-row
    semantics: ``hist[layer]`` is [n_slots, C, ctx]; all-zero == "no history", which is exactly
    what both a fresh cache (:337-346, :488-492) and set_to_zero (:234-241) mean for a causal conv.

    ``strict_reference_quirk=True`` reproduces the reference's late-joiner behaviour instead
    (SURVEY.md section 8(c) quirk 2C): ``cache.get`` returns None when ANY requested row has no entry
    for the layer (:204-207) and the callers then start EVERY row of the call from zero history
    (:337-346, :488-492); ``set`` creates the entries (:380, :547), ``set_to_zero`` keeps them (:234-241).
    """

    def __init__(self, n_slots: int, strict_reference_quirk: bool = False):
        self.n_slots = n_slots
        self.hist: Dict[str, torch.Tensor] = {}
        self.strict_reference_quirk = strict_reference_quirk
        self.present: Dict[str, set] = {}

    def get(self, layer: str, idx: torch.Tensor, C: int, ctx: int, like: torch.Tensor) -> torch.Tensor:
        if layer not in self.hist:
            self.hist[layer] = torch.zeros(self.n_slots, C, ctx, dtype=like.dtype)
        if self.strict_reference_quirk:
            have = self.present.setdefault(layer, set())
            if any(int(i) not in have for i in idx.tolist()):
                return torch.zeros(len(idx), C, ctx, dtype=like.dtype)      # get() -> None -> zeros for all rows
        return self.hist[layer][idx]

    def set(self, layer: str, idx: torch.Tensor, value: torch.Tensor) -> None:
        self.hist[layer][idx] = value
        if self.strict_reference_quirk:
            self.present.setdefault(layer, set()).update(int(i) for i in idx.tolist())

    def set_to_zero(self, idx: torch.Tensor) -> None:
        for v in self.hist.values():
            v[idx] = 0


def _sconv1d(x, w, b, groups, name, state, idx):
    """SConv1d, causal, stride 1, pad_mode='constant' (:258-418).  Non-streaming: left-pad k-1
    zeros (:398-401).  Streaming: prepend the cached last k-1 input samples (:335-380)."""
    k = w.shape[-1]
    ctx = k - 1                                             # context_size (:282), stride 1, dilation 1
    if state is None:
        xin = F.pad(x, (ctx, 0))
    else:
        h = state.get(name, idx, x.shape[1], ctx, x)
        xin = torch.cat([h, x], dim=2)
        state.set(name, idx, xin[:, :, -ctx:])
    return F.conv1d(xin, w, b, groups=groups)


def _sconvtr1d(x, w, b, stride, name, state, idx):
    """SConvTranspose1d, causal, trim_right_ratio=1 (:421-576): full conv_transpose1d, drop the
    last k - stride samples; streaming re-runs over up to k-1 cached input frames and returns
    the last T*stride samples (:486-547)."""
    k = w.shape[-1]
    pad_total = k - stride
    T = x.shape[2]
    if state is None:
        y = F.conv_transpose1d(x, w, b, stride=stride)
        return y[..., : y.shape[-1] - pad_total]
    ctx = k - 1                                             # context_size (:447)
    h = state.get(name, idx, x.shape[1], ctx, x)
    xin = torch.cat([h, x], dim=2)
    y = F.conv_transpose1d(xin, w, b, stride=stride)
    y = y[..., : y.shape[-1] - pad_total]
    state.set(name, idx, xin[:, :, -ctx:])
    return y[..., -T * stride:]


def decoder_forward(W: Dict[str, torch.Tensor], latents: torch.Tensor, depths, ratios,
                    eps: float = 1e-5, state: Optional[DecoderState] = None,
                    idx: Optional[torch.Tensor] = None, emulate_bf16: bool = False) -> torch.Tensor:
    """TokenizerDecoder.forward (:914-951): latents [n, vae, T] -> audio [n, 1, hop*T].

    state/idx given => streaming (use_cache=True) for cache rows idx; else zero history.
    emulate_bf16 rounds weights, the residual stream after every layer, the normalised
    conv/FFN inputs and the GELU output to bf16 (the CUDA pipeline's storage points).
    """
    r = _bf if emulate_bf16 else _ident
    x = r(latents.float())
    x = r(_sconv1d(x, r(W["upsample_layers.0.0.conv.conv.weight"]), W["upsample_layers.0.0.conv.conv.bias"],
                   1, "stem", state, idx))
    for i in range(len(depths)):
        if i > 0:
            p = f"upsample_layers.{i}.0.convtr.convtr."
            x = r(_sconvtr1d(x, r(W[p + "weight"]), W[p + "bias"], ratios[i - 1], f"up{i}", state, idx))
        for j in range(depths[i]):                          # inlined Block1D (:924-944)
            p = f"stages.{i}.{j}."
            C = x.shape[1]
            xn = r(conv_rmsnorm(x, W[p + "norm.weight"], eps))
            y = _sconv1d(xn, W[p + "mixer.conv.conv.conv.weight"], W[p + "mixer.conv.conv.conv.bias"],
                         C, f"dw{i}.{j}", state, idx)
            x = r(x + y * W[p + "gamma"].unsqueeze(-1))
            yn = r(conv_rmsnorm(x, W[p + "ffn_norm.weight"], eps)).permute(0, 2, 1)
            hmid = r(F.gelu(F.linear(yn, r(W[p + "ffn.linear1.weight"]), W[p + "ffn.linear1.bias"])))
            y = F.linear(hmid, r(W[p + "ffn.linear2.weight"]), W[p + "ffn.linear2.bias"]).permute(0, 2, 1)
            x = r(x + y * W[p + "ffn_gamma"].unsqueeze(-1))
    # self.norm is Identity (disable_last_norm, :908-911); head conv (:950)
    return _sconv1d(x, W["head.conv.conv.weight"], W["head.conv.conv.bias"], 1, "head", state, idx)


def decode(W, latents, depths, ratios, vae_dim, eps=1e-5, state=None, idx=None, emulate_bf16=False,
           scale: float = 1.0, bias: float = 0.0):
    """VibeVoiceAcousticTokenizerModel.decode (:1099-1108) preceded by the latent
    de-normalisation of modeling_vibevoice_inference.py:651 (z / scale - bias)."""
    if latents.shape[1] != vae_dim:
        latents = latents.permute(0, 2, 1)
    latents = latents.float() / scale - bias
    return decoder_forward(W, latents, depths, ratios, eps, state, idx, emulate_bf16)


# --------------------------------------------------------------------------------------
# Semantic tokenizer encoder and SpeechConnector (rank 2)
# --------------------------------------------------------------------------------------


def _sconv1d_strided(x, w, b, stride, name, state, idx):
    """SConv1d, causal, k = 2*stride (the encoder's downsample layers, :738-745).  context_size =
    padding_total = (k-1) - (stride-1) = stride (:282-285).  Non-streaming: left-pad `stride` zeros and
    right-pad the "extra padding" that completes the last window (get_extra_padding_for_conv1d, :127-133,
    :393-401: n_frames = (L - k + pad)/stride + 1, ideal = (ceil(n_frames)-1)*stride + k - pad, i.e. up to the
    next multiple of the stride).  Streaming: prepend the cached last `stride` input samples (:335-380)."""
    k = w.shape[-1]
    assert k == 2 * stride, (k, stride)
    ctx = k - stride
    if state is None:
        L = x.shape[2]
        n_frames = (L - k + ctx) / stride + 1
        ideal = (math.ceil(n_frames) - 1) * stride + (k - ctx)
        xin = F.pad(x, (ctx, ideal - L))
    else:
        assert x.shape[2] % stride == 0, (stride, x.shape)
        h = state.get(name, idx, x.shape[1], ctx, x)
        xin = torch.cat([h, x], dim=2)
        state.set(name, idx, xin[:, :, -ctx:])
    return F.conv1d(xin, w, b, stride=stride)


def encoder_forward(W: Dict[str, torch.Tensor], audio: torch.Tensor, depths, ratios,
                    eps: float = 1e-5, state: Optional[DecoderState] = None,
                    idx: Optional[torch.Tensor] = None, emulate_bf16: bool = False) -> torch.Tensor:
    """TokenizerEncoder.forward (:774-813): audio [n, 1, L] -> [n, dim, ceil(L/hop)] (streaming: L = hop*T).
    `ratios` in the order the
    encoder applies them (TokenizerEncoder reverses config.ratios, :701).  Block1D as in the decoder
    (inlined copy :785-806); self.norm is Identity (disable_last_norm, :767-770)."""
    r = _bf if emulate_bf16 else _ident
    x = r(audio.float())
    x = r(_sconv1d(x, W["downsample_layers.0.0.conv.conv.weight"], W["downsample_layers.0.0.conv.conv.bias"],
                   1, "stem", state, idx))
    for i in range(len(depths)):
        if i > 0:
            p = f"downsample_layers.{i}.0.conv.conv."
            x = r(_sconv1d_strided(x, r(W[p + "weight"]), W[p + "bias"], ratios[i - 1], f"down{i}", state, idx))
        for j in range(depths[i]):
            p = f"stages.{i}.{j}."
            C = x.shape[1]
            xn = r(conv_rmsnorm(x, W[p + "norm.weight"], eps))
            y = _sconv1d(xn, W[p + "mixer.conv.conv.conv.weight"], W[p + "mixer.conv.conv.conv.bias"],
                         C, f"dw{i}.{j}", state, idx)
            x = r(x + y * W[p + "gamma"].unsqueeze(-1))
            yn = r(conv_rmsnorm(x, W[p + "ffn_norm.weight"], eps)).permute(0, 2, 1)
            hmid = r(F.gelu(F.linear(yn, r(W[p + "ffn.linear1.weight"]), W[p + "ffn.linear1.bias"])))
            y = F.linear(hmid, r(W[p + "ffn.linear2.weight"]), W[p + "ffn.linear2.bias"]).permute(0, 2, 1)
            x = r(x + y * W[p + "ffn_gamma"].unsqueeze(-1))
    return _sconv1d(x, r(W["head.conv.conv.weight"]), W["head.conv.conv.bias"], 1, "head", state, idx)


def encode(W, audio, depths, ratios, eps=1e-5, state=None, idx=None, emulate_bf16=False) -> torch.Tensor:
    """VibeVoiceSemanticTokenizerModel.encode(...).mean (:1167-1170) and VibeVoiceAcousticTokenizerModel.encode(...).mean
    (:1081-1085): [n, 1, L] -> [n, ceil(L/hop), dim]."""
    if audio.dim() == 2:
        audio = audio.unsqueeze(1)
    return encoder_forward(W, audio, depths, ratios, eps, state, idx, emulate_bf16).permute(0, 2, 1)


def speech_connector(W: Dict[str, torch.Tensor], features: torch.Tensor, eps: float = 1e-6,
                     emulate_bf16: bool = False) -> torch.Tensor:
    """SpeechConnector.forward (modeling_vibevoice.py:58-69): fc1 -> LlamaRMSNorm(eps=1e-6) -> fc2.
    LlamaRMSNorm normalises in fp32, casts back, then multiplies by the weight."""
    r = _bf if emulate_bf16 else _ident
    x = F.linear(r(features.float()), r(W["fc1.weight"]), W["fc1.bias"])
    x = r(rmsnorm(x, W["norm.weight"], eps))
    return F.linear(x, r(W["fc2.weight"]), W["fc2.bias"])


def sample_latents(mean: torch.Tensor, std: float, dist_type: str, noise: Optional[torch.Tensor] = None,
                   row_noise: Optional[torch.Tensor] = None):
    """VibeVoiceTokenizerEncoderOutput.sample (:1018-1046) with the random draws passed in:
    'fix': mean + std * noise; 'gaussian': s = row_noise * (std / 0.8) broadcast over [T, dim], mean + s * noise;
    anything else: the mean.  Computed in mean's dtype, op by op, like the eager reference.  Returns (x, std used)."""
    if dist_type == "fix":
        return mean + std * noise, std
    if dist_type == "gaussian":
        s = row_noise.to(mean.dtype) * (std / 0.8)
        while s.dim() < mean.dim():
            s = s.unsqueeze(-1)
        return mean + s * noise, s
    return mean, std


def acoustic_features(latents: torch.Tensor, bias_factor, scaling_factor) -> torch.Tensor:
    """(latents + speech_bias_factor) * speech_scaling_factor (modeling_vibevoice_inference.py:159, :170); the
    factors are 0-dim buffers, so the result keeps the latents' dtype."""
    b = torch.as_tensor(bias_factor)
    s = torch.as_tensor(scaling_factor)
    return (latents + b) * s
