// vv_head.cu — diffusion head + DPM-Solver++ sampling loop.
//
// Reference: VibeVoiceDiffusionHead.forward (modular_vibevoice_diffusion_head.py:254-280),
// sample_speech_tokens (modeling_vibevoice_inference.py:712-725), DPMSolverMultistepScheduler
// (vibevoice/schedule/dpm_solver.py:321-423, 528-594, 627-817, 935-1022).
//
// Data layout in HBM (all row-major, M = 2n rows: cond rows 0..n-1, uncond rows n..2n-1):
//   x      f32  [M,H]    residual stream (kept fp32; the reference rounds it to bf16 each op)
//   y      bf16 [M,H]    modulated RMSNorm output = A operand of the gate|up GEMM
//   h      bf16 [M,F]    silu(gate)*up            = A operand of the down GEMM
//   mod    bf16 [M,NADA] all adaLN outputs of one step: per layer (shift|scale|gate), then the
//                        final layer's (shift|scale); c = cond_proj(cond)+t_emb does not depend
//                        on x, so one GEMM per step produces every layer's modulation
//   W_gu   bf16 [2F,H]   gate/up rows interleaved so the SwiGLU pair sits in adjacent
//                        accumulator columns (fused in the GEMM epilogue)
// Small M (streaming) runs the down / final projections as split-K partials that the next
// row-wise kernel reduces (deterministic, no atomics); large M fuses the gated residual into
// the GEMM epilogue.
#include <math.h>

#include <map>
#include <vector>

#include "../../include/vvb200.h"
#include "vv_common.cuh"
#include "vv_chain.cuh"
#include "vv_gemm.cuh"

namespace {

// ------------------------------- packing kernels ------------------------------------
__global__ void k_f32_to_bf16(const float* __restrict__ in, bf16* __restrict__ out, size_t n) {
  size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (; i < n; i += stride) out[i] = __float2bfloat16(in[i]);
}
__global__ void k_bf16_to_f32(const bf16* __restrict__ in, float* __restrict__ out, size_t n) {
  size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (; i < n; i += stride) out[i] = __bfloat162float(in[i]);
}
// out[2j] = gate[j], out[2j+1] = up[j]   (rows of length K)
template <typename T>
__global__ void k_interleave_rows(const T* __restrict__ gate, const T* __restrict__ up,
                                  bf16* __restrict__ out, int F, int K) {
  const size_t total = static_cast<size_t>(2) * F * K;
  size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (; i < total; i += stride) {
    const int r = static_cast<int>(i / K), k = static_cast<int>(i % K);
    const T* src = (r & 1) ? up : gate;
    out[i] = __float2bfloat16(static_cast<float>(src[static_cast<size_t>(r >> 1) * K + k]));
  }
}

// ------------------------------- head kernels ---------------------------------------
// TimestepEmbedder.timestep_embedding (modular_vibevoice_diffusion_head.py:66-88): cat[cos, sin]
__global__ void k_sinusoid(const float* __restrict__ t, bf16* __restrict__ out, int M, int dim) {
  const int half = dim / 2;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= M * half) return;
  const int m = i / half, k = i % half;
  const float freq = expf(-logf(10000.0f) * static_cast<float>(k) / static_cast<float>(half));
  const float a = t[m] * freq;
  out[static_cast<size_t>(m) * dim + k] = __float2bfloat16(cosf(a));
  out[static_cast<size_t>(m) * dim + half + k] = __float2bfloat16(sinf(a));
}

// sc = bf16(silu(cproj + temb[row]))  — the SiLU in front of every adaLN Linear (:152-156, :178-182)
//   row r = s * M + m  (s < S):  temb row = s * temb_step_stride + m * temb_row_stride
// sample(): S solver steps share one t-embedding per step (temb_row_stride = 0); forward(): S = 1
// and every row has its own (temb_step_stride = 0).
__global__ void k_silu_add(const float* __restrict__ cproj, const float* __restrict__ temb,
                           int temb_row_stride, int temb_step_stride, bf16* __restrict__ out, int M, int S,
                           int H) {
  griddep_wait();
  griddep_launch();
  const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;   // one float4 per thread
  const int per_row = H / 4;
  if (i >= static_cast<size_t>(M) * S * per_row) return;
  const size_t r = i / per_row;
  const int c = static_cast<int>(i % per_row) * 4;
  const int m = static_cast<int>(r % M), st = static_cast<int>(r / M);
  const float4 a = *reinterpret_cast<const float4*>(cproj + static_cast<size_t>(m) * H + c);
  const float4 b = *reinterpret_cast<const float4*>(temb + static_cast<size_t>(st) * temb_step_stride +
                                                    static_cast<size_t>(m) * temb_row_stride + c);
  uint2 o;
  o.x = pack_bf16x2(silu_f(a.x + b.x), silu_f(a.y + b.y));
  o.y = pack_bf16x2(silu_f(a.z + b.z), silu_f(a.w + b.w));
  *reinterpret_cast<uint2*>(out + r * H + c) = o;
}

// One block per row.  Optionally folds the previous layer's split-K partials and gate into the
// residual (x += gate * sum_z partial_z), then y = bf16(rmsnorm(x) * w * (1 + scale) + shift).
//   RMSNorm (:31-38), modulate (:43-45), HeadLayer.forward (:158-161), FinalLayer.forward (:184-188)
constexpr int NM_THREADS = 256;
constexpr int NM_MAX_VEC = 4;   // H <= 4 * 4 * 256 = 4096
__global__ void __launch_bounds__(NM_THREADS)
k_norm_mod(float* __restrict__ x, const float* __restrict__ partial, int nsplit, int M,
           const bf16* __restrict__ prev_gate, const float* __restrict__ w,
           const bf16* __restrict__ shift, const bf16* __restrict__ scale, int ld_mod,
           bf16* __restrict__ y, int H, float eps) {
  griddep_wait();
  griddep_launch();
  const int m = blockIdx.x;
  const int nvec = H / 4;
  float4 v[NM_MAX_VEC];
  float ss = 0.f;
  float* xr = x + static_cast<size_t>(m) * H;
#pragma unroll
  for (int j = 0; j < NM_MAX_VEC; ++j) {
    const int i = threadIdx.x + j * NM_THREADS;
    if (i < nvec) {
      float4 a = reinterpret_cast<const float4*>(xr)[i];
      if (nsplit > 0) {
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int z0 = 0; z0 < nsplit; z0 += 8) {     // 8 independent loads in flight, then the adds
          float4 q[8];
#pragma unroll
          for (int u = 0; u < 8; ++u)
            q[u] = (z0 + u < nsplit)
                       ? reinterpret_cast<const float4*>(partial + (static_cast<size_t>(z0 + u) * M + m) * H)[i]
                       : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int u = 0; u < 8; ++u) { s.x += q[u].x; s.y += q[u].y; s.z += q[u].z; s.w += q[u].w; }
        }
        const uint2 gv = *reinterpret_cast<const uint2*>(prev_gate + static_cast<size_t>(m) * ld_mod + i * 4);
        const float2 g0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gv.x));
        const float2 g1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gv.y));
        a.x += g0.x * s.x; a.y += g0.y * s.y; a.z += g1.x * s.z; a.w += g1.y * s.w;
        reinterpret_cast<float4*>(xr)[i] = a;
      }
      v[j] = a;
      ss += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w;
    }
  }
  __shared__ float red[NM_THREADS / 32];
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int i = 0; i < NM_THREADS / 32; ++i) tot += red[i];
  const float rms = rsqrtf(tot / static_cast<float>(H) + eps);
#pragma unroll
  for (int j = 0; j < NM_MAX_VEC; ++j) {
    const int i = threadIdx.x + j * NM_THREADS;
    if (i < nvec) {
      float4 wv = make_float4(1.f, 1.f, 1.f, 1.f);
      if (w != nullptr) wv = __ldg(reinterpret_cast<const float4*>(w) + i);
      const uint2 sh = *reinterpret_cast<const uint2*>(shift + static_cast<size_t>(m) * ld_mod + i * 4);
      const uint2 sc = *reinterpret_cast<const uint2*>(scale + static_cast<size_t>(m) * ld_mod + i * 4);
      const float2 sh0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&sh.x));
      const float2 sh1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&sh.y));
      const float2 sc0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&sc.x));
      const float2 sc1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&sc.y));
      const float4 a = v[j];
      uint2 o;
      o.x = pack_bf16x2(a.x * rms * wv.x * (1.f + sc0.x) + sh0.x, a.y * rms * wv.y * (1.f + sc0.y) + sh0.y);
      o.y = pack_bf16x2(a.z * rms * wv.z * (1.f + sc1.x) + sh1.x, a.w * rms * wv.w * (1.f + sc1.y) + sh1.y);
      *reinterpret_cast<uint2*>(y + static_cast<size_t>(m) * H + i * 4) = o;
    }
  }
}

// ------------------------------- solver kernels -------------------------------------
struct StepCoef {      // one DPM-Solver++ step (see oracle/vv_oracle.py:solver_coefficients)
  float alpha_s, sigma_s;   // x0 = alpha_s * x - sigma_s * v          (dpm_solver.py:581-584)
  float cx, c0, c1, inv_r0; // x' = cx*x + c0*x0 + c1*(x0 - x0_prev)*inv_r0   (:669-677, :752-764)
  float cn;                 // + cn * noise for sde-dpmsolver++         (:680-686, :785-793)
  int order;
};

template <typename T>
__global__ void k_to_bf16_rows(const T* __restrict__ in, bf16* __restrict__ out, size_t n) {
  size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  if (i < n) out[i] = __float2bfloat16(static_cast<float>(in[i]));
}

// zstate = noise[:n] (speech[:n], modeling_vibevoice_inference.py:716-718); zb = bf16([z; z]) (:719)
__global__ void k_init_state(const float* __restrict__ noise, float* __restrict__ zstate,
                             bf16* __restrict__ zb, int n, int L) {
  griddep_wait();
  griddep_launch();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * L) return;
  const float v = noise[i];
  zstate[i] = v;
  const bf16 b = __float2bfloat16(v);
  zb[i] = b;
  zb[static_cast<size_t>(n) * L + i] = b;
}

// CFG combine + v->x0 + DPM-Solver++ update, one thread per latent element.
//   v: [nsplit][2n][L] split-K partials of the final Linear (nsplit >= 1)
//   (modeling_vibevoice_inference.py:721-724; dpm_solver.py:987-1014)
__global__ void k_cfg_update(const float* __restrict__ v, int nsplit, float* __restrict__ zstate,
                             float* __restrict__ prev_x0, bf16* __restrict__ zb,
                             const StepCoef* __restrict__ coefs, int step, float cfg,
                             const float* __restrict__ sde_noise, void* __restrict__ out,
                             int out_dtype, int n, int L) {
  griddep_wait();
  griddep_launch();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * L) return;
  const StepCoef k = coefs[step];
  const size_t MN = static_cast<size_t>(2) * n * L;
  float c = 0.f, u = 0.f;
  for (int z = 0; z < nsplit; ++z) {
    c += v[z * MN + i];
    u += v[z * MN + static_cast<size_t>(n) * L + i];
  }
  const float veps = u + cfg * (c - u);
  const float x = zstate[i];
  const float x0 = k.alpha_s * x - k.sigma_s * veps;
  float xn = k.cx * x + k.c0 * x0;
  if (k.order == 2) xn += k.c1 * (k.inv_r0 * (x0 - prev_x0[i]));
  if (k.cn != 0.f && sde_noise != nullptr) xn += k.cn * sde_noise[i];
  prev_x0[i] = x0;
  zstate[i] = xn;
  const bf16 b = __float2bfloat16(xn);
  zb[i] = b;
  zb[static_cast<size_t>(n) * L + i] = b;
  if (out != nullptr) {
    if (out_dtype == VV_DTYPE_BF16) reinterpret_cast<bf16*>(out)[i] = b;
    else reinterpret_cast<float*>(out)[i] = xn;
  }
}

// forward(): sum split-K partials of the final Linear and store in the caller's dtype.
__global__ void k_store_out(const float* __restrict__ v, int nsplit, size_t MN, void* __restrict__ out,
                            int out_dtype) {
  const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  if (i >= MN) return;
  float s = 0.f;
  for (int z = 0; z < nsplit; ++z) s += v[z * MN + i];
  if (out_dtype == VV_DTYPE_BF16) reinterpret_cast<bf16*>(out)[i] = __float2bfloat16(s);
  else reinterpret_cast<float*>(out)[i] = s;
}

inline dim3 grid1d(size_t n, int block) { return dim3(static_cast<unsigned>((n + block - 1) / block)); }

}  // namespace

// ------------------------------------------------------------------------------------
// Handle
// ------------------------------------------------------------------------------------
struct VVHead {
  VVHeadDesc d;
  int H, F, L, NL, NADA;
  // packed weights
  DevBuf w_in, w_cond, w_t0, w_t1, w_ada, w_out, norm_w;
  std::vector<DevBuf> w_gu, w_down;
  // schedule
  int S = 0, algo = 0;
  DevBuf coefs, temb_table;     // [S] StepCoef ; [S,H] f32
  std::vector<StepCoef> coefs_host;
  // scratch (sized for d.max_rows)
  DevBuf zb, condb, cproj, temb_rows, tfreq, th, sc, mod, x, y, h, partial, vpart;
  DevBuf zstate, prev_x0, tvals;
  // persistent solver loop (vv_chain.cu): one program per batch size (a streaming server steps a
  // varying subset of rows), all invalidated when the schedule or the scratch buffers change
  static constexpr int MAX_CHAINS = 64;   // one per batch size a 64-row generate() can reach: no eviction mid-generation
  std::map<int, ChainProgram*> chains;   // n -> program; nullptr: this n cannot run as a chain (falls back)
  int chain_S = -1;
  const void* chain_mod = nullptr;
  void drop_chains() {
    for (auto& kv : chains) if (kv.second) chain_destroy(kv.second);
    chains.clear();
  }
  ~VVHead() { drop_chains(); }
};

namespace {

int upload_bf16(DevBuf& dst, const void* src, size_t n, int src_dtype, cudaStream_t st) {
  VV_TRY(dst.alloc(n * sizeof(bf16)));
  if (src_dtype == VV_DTYPE_BF16) {
    VV_CHECK_CUDA(cudaMemcpyAsync(dst.p, src, n * sizeof(bf16), cudaMemcpyDeviceToDevice, st));
  } else {
    k_f32_to_bf16<<<1024, 256, 0, st>>>(static_cast<const float*>(src), dst.as<bf16>(), n);
    VV_COUNT_LAUNCH();
    VV_CHECK_CUDA(cudaGetLastError());
  }
  return VV_OK;
}
int upload_f32(float* dst, const void* src, size_t n, int src_dtype, cudaStream_t st) {
  if (src_dtype == VV_DTYPE_F32) {
    VV_CHECK_CUDA(cudaMemcpyAsync(dst, src, n * sizeof(float), cudaMemcpyDeviceToDevice, st));
  } else {
    k_bf16_to_f32<<<256, 256, 0, st>>>(static_cast<const bf16*>(src), dst, n);
    VV_COUNT_LAUNCH();
    VV_CHECK_CUDA(cudaGetLastError());
  }
  return VV_OK;
}

int head_gemm(const bf16* A, int lda, const bf16* W, int ldw, int M, int N, int K, int mode, void* out,
              int ldo, const void* resid, int ldr, const bf16* gate, int ldg, int* splits_out,
              cudaStream_t st) {
  GemmArgs a;
  a.A = A; a.lda = lda; a.B = W; a.ldb = ldw; a.M = M; a.N = N; a.K = K; a.mode = mode;
  a.out = out; a.ldo = ldo; a.resid = resid; a.ldr = ldr; a.gate = gate; a.ldg = ldg;
  a.b_streaming = (M <= 256);
  gemm_plan(a, vv_num_sms());
  if (splits_out) *splits_out = a.splits;
  return gemm_launch(a, st);
}

constexpr int SMALL_M = 256;      // at or below: split-K partial path for the narrow-N GEMMs
constexpr int MAX_SPLITS = 32;
constexpr size_t MOD_TABLE_BUDGET = static_cast<size_t>(512) << 20;   // all-steps adaLN table (55 MB at n=64, S=10)

// adaLN modulation of every layer for S "steps" of M rows in ONE GEMM: the input silu(c),
// c = cond_proj(cond) + t_emb, does not depend on x, so all solver steps of a frame are batched
// (M*S rows against the 66 MB of adaLN weights, read once per frame instead of once per step).
//   mod[(s*M + m), :] = [layer0 (shift|scale|gate) | ... | final (shift|scale)]      (:159, :185)
int head_modulation(VVHead* h, int M, int S, const float* temb, int temb_row_stride, int temb_step_stride,
                    cudaStream_t st) {
  const int H = h->H, NADA = h->NADA;
  const size_t rows = static_cast<size_t>(M) * S;
  VV_TRY(h->sc.ensure(rows * H * sizeof(bf16)));
  VV_TRY(h->mod.ensure(rows * NADA * sizeof(bf16)));
  VV_CHECK_CUDA(vv_launch(k_silu_add, grid1d(rows * H / 4, 256), dim3(256), 0, st, 1, true, h->cproj.as<float>(),
                          temb, temb_row_stride, temb_step_stride, h->sc.as<bf16>(), M, S, H));
  VV_TRY(head_gemm(h->sc.as<bf16>(), H, h->w_ada.as<bf16>(), H, static_cast<int>(rows), NADA, H, GM_STORE_BF16,
                   h->mod.p, NADA, nullptr, 0, nullptr, 0, nullptr, st));
  return VV_OK;
}

// Core of VibeVoiceDiffusionHead.forward given zb (bf16 [M,L]) and the rows' modulation `mod`
// ([M,NADA] bf16).  Leaves split-K partials of the final Linear in h->vpart ([*nsplit_v][M][L]).
int head_core(VVHead* h, int M, const bf16* mod, int* nsplit_v, cudaStream_t st) {
  const int H = h->H, F = h->F, L = h->L, NADA = h->NADA;
  // x = noisy_images_proj(z) (:271)
  VV_TRY(head_gemm(h->zb.as<bf16>(), L, h->w_in.as<bf16>(), L, M, H, L, GM_STORE_F32, h->x.p, H,
                   nullptr, 0, nullptr, 0, nullptr, st));
  int pending = 0;                 // split-K partials of the previous down-projection
  const bf16* pending_gate = nullptr;
  for (int l = 0; l < h->NL; ++l) {
    const bf16* shift = mod + static_cast<size_t>(l) * 3 * H;
    const bf16* scale = shift + H;
    const bf16* gate = shift + 2 * H;
    VV_CHECK_CUDA(vv_launch(k_norm_mod, dim3(M), dim3(NM_THREADS), 0, st, 1, true, h->x.as<float>(),
                            h->partial.as<float>(), pending, M, pending_gate,
                            h->norm_w.as<float>() + static_cast<size_t>(l) * H, shift, scale, NADA,
                            h->y.as<bf16>(), H, h->d.eps));
    VV_TRY(head_gemm(h->y.as<bf16>(), H, h->w_gu[l].as<bf16>(), H, M, 2 * F, H, GM_SWIGLU_BF16,
                     h->h.p, F, nullptr, 0, nullptr, 0, nullptr, st));
    if (M <= SMALL_M) {
      VV_TRY(head_gemm(h->h.as<bf16>(), F, h->w_down[l].as<bf16>(), F, M, H, F, GM_PARTIAL_F32,
                       h->partial.p, H, nullptr, 0, nullptr, 0, &pending, st));
      pending_gate = gate;
    } else {
      VV_TRY(head_gemm(h->h.as<bf16>(), F, h->w_down[l].as<bf16>(), F, M, H, F, GM_RESID_GATE_F32,
                       h->x.p, H, h->x.p, H, gate, NADA, nullptr, st));
      pending = 0;
    }
  }
  const bf16* fshift = mod + static_cast<size_t>(h->NL) * 3 * H;
  VV_CHECK_CUDA(vv_launch(k_norm_mod, dim3(M), dim3(NM_THREADS), 0, st, 1, true, h->x.as<float>(),
                          h->partial.as<float>(), pending, M, pending_gate, static_cast<const float*>(nullptr),
                          fshift, fshift + H, NADA, h->y.as<bf16>(), H, h->d.eps));
  VV_TRY(head_gemm(h->y.as<bf16>(), H, h->w_out.as<bf16>(), H, M, L, H, GM_PARTIAL_F32, h->vpart.p, L,
                   nullptr, 0, nullptr, 0, nsplit_v, st));
  return VV_OK;
}

// The S-step solver loop as ONE persistent launch: the same op sequence as head_core + k_cfg_update,
// expressed as a phase program for the chain engine.  Needs the all-steps modulation table.
int head_build_chain(VVHead* h, int n, ChainProgram* cp) {
  const int H = h->H, F = h->F, L = h->L, NADA = h->NADA, M = 2 * n, S = h->S;
  chain_clear(cp);
  const int sms = vv_num_sms();
  auto splits_for = [&](int N, int K) {
    const int tiles = vv_ceil_div(N, 64), kb = vv_ceil_div(K, 64);
    int s = sms / tiles;
    if (s < 1) s = 1;
    const int max_s = kb / 4 > 0 ? kb / 4 : 1;
    if (s > max_s) s = max_s;
    if (s > MAX_SPLITS) s = MAX_SPLITS;
    while (s > 1 && vv_ceil_div(kb, s) * (s - 1) >= kb) --s;
    return s;
  };
  const int sp_down = splits_for(H, F), sp_out = splits_for(L, H);
  auto gemm = [&](const bf16* A, int lda, const bf16* W, int ldw, int N, int K, int mode, void* out, int ldo, int splits,
                  bool keep = false, int rowmajor = 0) {
    GemmArgs a;
    a.A = A; a.lda = lda; a.B = W; a.ldb = ldw; a.M = M; a.N = N; a.K = K; a.mode = mode;
    a.out = out; a.ldo = ldo; a.splits = splits;
    a.b_streaming = keep ? 2 : 1;
    return chain_add_gemm(cp, a, rowmajor);
  };
  // the down projection's split-K partials lie [M][splits][H]: the row phase that sums them fetches a row's with one bulk copy
  const int part_rm = vv_option_chain_part_rowmajor() ? 1 : 0;
  // the S solver steps re-read the same weights (170 MB per step against 126 MB of L2): the first `keep_layers` layers'
  // matrices are loaded evict-last so that they may survive from one step to the next
  const int keep_layers = vv_option_chain_keep_layers();
  const bool fuse_tail = (L == 64 && H % 128 == 0) && vv_option_head_fuse_tail();   // step boundary as one row-pair phase (ChainTail)
  for (int s = 0; s < S; ++s) {
    const bf16* mod = h->mod.as<bf16>() + static_cast<size_t>(s) * M * NADA;
    const bool head_of_step = (s == 0 || !fuse_tail);  // otherwise the previous tail already produced x and y
    if (head_of_step) VV_TRY(gemm(h->zb.as<bf16>(), L, h->w_in.as<bf16>(), L, H, L, GM_STORE_F32, h->x.p, H, 1));
    int pending = 0;
    const bf16* pending_gate = nullptr;
    for (int l = 0; l <= h->NL; ++l) {
      const bool fin = (l == h->NL);
      const bf16* shift = mod + static_cast<size_t>(l) * 3 * H;
      if (fin && fuse_tail) break;
      if (l > 0 || head_of_step) {
        ChainNormMod nm;
        nm.x = h->x.as<float>(); nm.partial = h->partial.as<float>(); nm.nsplit = pending; nm.M = M; nm.rowmajor = part_rm;
        nm.prev_gate = pending_gate;
        nm.w = fin ? nullptr : h->norm_w.as<float>() + static_cast<size_t>(l) * H;
        nm.shift = shift; nm.scale = shift + H; nm.ld_mod = NADA;
        nm.y = h->y.as<bf16>(); nm.H = H; nm.eps = h->d.eps;
        VV_TRY(chain_add_norm_mod(cp, nm));
      }
      if (fin) break;
      VV_TRY(gemm(h->y.as<bf16>(), H, h->w_gu[l].as<bf16>(), H, 2 * F, H, GM_SWIGLU_BF16, h->h.p, F, 1, l < keep_layers));
      VV_TRY(gemm(h->h.as<bf16>(), F, h->w_down[l].as<bf16>(), F, H, F, GM_PARTIAL_F32, h->partial.p, H, sp_down, l < keep_layers, part_rm));
      pending = sp_down;
      pending_gate = shift + 2 * H;
    }
    const StepCoef& k = h->coefs_host[s];
    const bool last = (s == S - 1);
    if (fuse_tail) {
      ChainTail t{};
      t.x = h->x.as<float>(); t.partial = h->partial.as<float>(); t.prev_gate = pending_gate;
      t.fshift = mod + static_cast<size_t>(h->NL) * 3 * H;
      t.w_out = h->w_out.as<bf16>(); t.zstate = h->zstate.as<float>(); t.prev_x0 = h->prev_x0.as<float>();
      t.zb = h->zb.as<bf16>(); t.w_in = h->w_in.as<bf16>(); t.norm_w0 = h->norm_w.as<float>();
      t.shift0 = last ? nullptr : h->mod.as<bf16>() + static_cast<size_t>(s + 1) * M * NADA;
      t.y = h->y.as<bf16>();
      t.rowmajor = part_rm;
      t.nsplit = pending; t.n = n; t.ld_mod = NADA; t.step = s; t.last = last ? 1 : 0; t.L = L; t.H = H;
      t.order = k.order; t.alpha_s = k.alpha_s; t.sigma_s = k.sigma_s; t.cx = k.cx; t.c0 = k.c0; t.c1 = k.c1;
      t.inv_r0 = k.inv_r0; t.cn = k.cn; t.eps = h->d.eps;
      VV_TRY(chain_add_tail(cp, t));
      continue;
    }
    VV_TRY(gemm(h->y.as<bf16>(), H, h->w_out.as<bf16>(), H, L, H, GM_PARTIAL_F32, h->vpart.p, L, sp_out));
    ChainCfg cf;
    cf.v = h->vpart.as<float>(); cf.nsplit = sp_out; cf.zstate = h->zstate.as<float>();
    cf.prev_x0 = h->prev_x0.as<float>(); cf.zb = h->zb.as<bf16>(); cf.step = s;
    cf.alpha_s = k.alpha_s; cf.sigma_s = k.sigma_s; cf.cx = k.cx; cf.c0 = k.c0; cf.c1 = k.c1;
    cf.inv_r0 = k.inv_r0; cf.cn = k.cn; cf.order = k.order;
    cf.last = last ? 1 : 0; cf.n = n; cf.L = L;
    VV_TRY(chain_add_cfg(cp, cf));
  }
  VV_TRY(chain_finalize(cp));
  return VV_OK;
}

int to_bf16(const void* src, bf16* dst, size_t n, int dtype, cudaStream_t st) {
  if (dtype == VV_DTYPE_BF16) {
    VV_CHECK_CUDA(cudaMemcpyAsync(dst, src, n * sizeof(bf16), cudaMemcpyDeviceToDevice, st));
  } else {
    k_to_bf16_rows<float><<<grid1d(n, 256), 256, 0, st>>>(static_cast<const float*>(src), dst, n);
    VV_COUNT_LAUNCH();
  }
  return VV_OK;
}

}  // namespace

// ------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------
extern "C" {

int vv_head_create(const VVHeadDesc* desc, const void* const* weights, int n_weights, VVHead** out) {
  VV_REQUIRE(desc && weights && out, "vv_head_create: null argument");
  const int H = desc->hidden, F = desc->ffn, L = desc->latent, NL = desc->layers;
  VV_REQUIRE(H > 0 && H % 64 == 0 && H <= 4096, "vv_head_create: hidden=%d must be a multiple of 64, <= 4096", H);
  VV_REQUIRE(F > 0 && F % 64 == 0, "vv_head_create: ffn=%d must be a multiple of 64", F);
  VV_REQUIRE(L > 0 && L % 8 == 0, "vv_head_create: latent=%d must be a multiple of 8", L);
  VV_REQUIRE(desc->freq_dim % 8 == 0 && desc->freq_dim > 0, "vv_head_create: freq_dim");
  VV_REQUIRE(NL >= 1 && NL <= 64, "vv_head_create: layers=%d", NL);
  VV_REQUIRE(n_weights == 4 + 5 * NL + 2, "vv_head_create: expected %d weights, got %d", 4 + 5 * NL + 2, n_weights);
  VV_REQUIRE(desc->max_rows >= 2, "vv_head_create: max_rows=%d", desc->max_rows);
  for (int i = 0; i < n_weights; ++i) VV_REQUIRE(weights[i] != nullptr, "vv_head_create: weight %d is null", i);
  VVHead* h = new VVHead();
  h->d = *desc; h->H = H; h->F = F; h->L = L; h->NL = NL;
  h->NADA = 3 * H * NL + 2 * H;
  const int wd = desc->weight_dtype;
  cudaStream_t st = 0;
  int rc = VV_OK;
  auto fail = [&](int code) { delete h; return code; };
#define HTRY(e) do { rc = (e); if (rc != VV_OK) return fail(rc); } while (0)
  HTRY(upload_bf16(h->w_in, weights[0], static_cast<size_t>(H) * L, wd, st));
  HTRY(upload_bf16(h->w_cond, weights[1], static_cast<size_t>(H) * H, wd, st));
  HTRY(upload_bf16(h->w_t0, weights[2], static_cast<size_t>(H) * desc->freq_dim, wd, st));
  HTRY(upload_bf16(h->w_t1, weights[3], static_cast<size_t>(H) * H, wd, st));
  HTRY(h->w_ada.alloc(static_cast<size_t>(h->NADA) * H * sizeof(bf16)));
  HTRY(h->norm_w.alloc(static_cast<size_t>(NL) * H * sizeof(float)));
  h->w_gu = std::vector<DevBuf>(NL);
  h->w_down = std::vector<DevBuf>(NL);
  for (int l = 0; l < NL; ++l) {
    const void* const* w = weights + 4 + 5 * l;
    HTRY(upload_f32(h->norm_w.as<float>() + static_cast<size_t>(l) * H, w[0], H, wd, st));
    bf16* ada_dst = h->w_ada.as<bf16>() + static_cast<size_t>(l) * 3 * H * H;
    const size_t n_ada = static_cast<size_t>(3) * H * H;
    if (wd == VV_DTYPE_BF16) {
      if (cudaMemcpyAsync(ada_dst, w[1], n_ada * 2, cudaMemcpyDeviceToDevice, st) != cudaSuccess) return fail(VV_ERR_CUDA);
    } else {
      k_f32_to_bf16<<<1024, 256, 0, st>>>(static_cast<const float*>(w[1]), ada_dst, n_ada);
    }
    HTRY(h->w_gu[l].alloc(static_cast<size_t>(2) * F * H * sizeof(bf16)));
    if (wd == VV_DTYPE_BF16)
      k_interleave_rows<bf16><<<1024, 256, 0, st>>>(static_cast<const bf16*>(w[2]), static_cast<const bf16*>(w[3]),
                                                    h->w_gu[l].as<bf16>(), F, H);
    else
      k_interleave_rows<float><<<1024, 256, 0, st>>>(static_cast<const float*>(w[2]), static_cast<const float*>(w[3]),
                                                     h->w_gu[l].as<bf16>(), F, H);
    HTRY(upload_bf16(h->w_down[l], w[4], static_cast<size_t>(H) * F, wd, st));
  }
  {
    const void* const* w = weights + 4 + 5 * NL;
    bf16* ada_dst = h->w_ada.as<bf16>() + static_cast<size_t>(NL) * 3 * H * H;
    const size_t n_ada = static_cast<size_t>(2) * H * H;
    if (wd == VV_DTYPE_BF16) {
      if (cudaMemcpyAsync(ada_dst, w[0], n_ada * 2, cudaMemcpyDeviceToDevice, st) != cudaSuccess) return fail(VV_ERR_CUDA);
    } else {
      k_f32_to_bf16<<<1024, 256, 0, st>>>(static_cast<const float*>(w[0]), ada_dst, n_ada);
    }
    HTRY(upload_bf16(h->w_out, w[1], static_cast<size_t>(L) * H, wd, st));
  }
  const size_t Mx = desc->max_rows;
  const size_t Ms = Mx < SMALL_M ? Mx : SMALL_M;
  HTRY(h->zb.alloc(Mx * L * 2));
  HTRY(h->condb.alloc(Mx * H * 2));
  HTRY(h->cproj.alloc(Mx * H * 4));
  HTRY(h->temb_rows.alloc(Mx * H * 4));
  HTRY(h->tfreq.alloc(Mx * desc->freq_dim * 2));
  HTRY(h->th.alloc(Mx * H * 2));
  HTRY(h->sc.alloc(Mx * H * 2));
  HTRY(h->mod.alloc(Mx * h->NADA * 2));
  HTRY(h->x.alloc(Mx * H * 4));
  HTRY(h->y.alloc(Mx * H * 2));
  HTRY(h->h.alloc(Mx * F * 2));
  HTRY(h->partial.alloc(static_cast<size_t>(MAX_SPLITS) * Ms * H * 4));
  HTRY(h->vpart.alloc(static_cast<size_t>(MAX_SPLITS) * Mx * L * 4));
  HTRY(h->zstate.alloc(Mx * L * 4));
  HTRY(h->prev_x0.alloc(Mx * L * 4));
  HTRY(h->tvals.alloc(Mx * 4));
#undef HTRY
  VV_COUNT_SYNC();
  cudaError_t e = cudaStreamSynchronize(st);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) {
    vv_set_error("vv_head_create: %s", cudaGetErrorString(e));
    return fail(VV_ERR_CUDA);
  }
  *out = h;
  return VV_OK;
}

void vv_head_destroy(VVHead* h) { delete h; }

}  // extern "C"

namespace {

// t [rows] (device f32) -> temb_out [rows,H] f32 : sinusoid -> Linear -> SiLU -> Linear (:90-93)
int t_embed_rows(VVHead* h, const float* t_dev, int rows, bf16* tfreq, bf16* th, float* temb_out,
                 cudaStream_t st) {
  const int H = h->H, FD = h->d.freq_dim;
  k_sinusoid<<<grid1d(static_cast<size_t>(rows) * FD / 2, 256), 256, 0, st>>>(t_dev, tfreq, rows, FD);
  VV_COUNT_LAUNCH();
  VV_TRY(head_gemm(tfreq, FD, h->w_t0.as<bf16>(), FD, rows, H, FD, GM_SILU_BF16, th, H, nullptr, 0,
                   nullptr, 0, nullptr, st));
  VV_TRY(head_gemm(th, H, h->w_t1.as<bf16>(), H, rows, H, H, GM_STORE_F32, temb_out, H, nullptr, 0,
                   nullptr, 0, nullptr, st));
  return VV_OK;
}

void alpha_sigma(float s, float* a, float* st) {   // _sigma_to_alpha_sigma_t (dpm_solver.py:483-487)
  *a = 1.0f / sqrtf(s * s + 1.0f);
  *st = s * (*a);
}

}  // namespace

extern "C" {

int vv_head_set_schedule(VVHead* h, int S, const float* t_values, const float* sigmas, int algo,
                         void* stream) {
  VV_REQUIRE(h && t_values && sigmas, "vv_head_set_schedule: null argument");
  VV_REQUIRE(S >= 1 && S <= 1000, "vv_head_set_schedule: S=%d", S);
  VV_REQUIRE(algo == VV_ALGO_DPMPP || algo == VV_ALGO_SDE_DPMPP, "vv_head_set_schedule: algo=%d", algo);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  std::vector<StepCoef> co(S);
  for (int i = 0; i < S; ++i) {
    const bool first = (i == 0), last = (i == S - 1);   // lower_order_final (dpm_solver.py:978-982)
    StepCoef c{};
    c.order = (first || last) ? 1 : 2;
    float a_s0, s_s0, a_t, s_t;
    alpha_sigma(sigmas[i], &a_s0, &s_s0);
    alpha_sigma(sigmas[i + 1], &a_t, &s_t);
    const float lam_t = logf(a_t) - logf(s_t);
    const float lam_s0 = logf(a_s0) - logf(s_s0);
    const float hh = lam_t - lam_s0;
    c.alpha_s = a_s0;
    c.sigma_s = s_s0;
    if (c.order == 2) {
      float a_s1, s_s1;
      alpha_sigma(sigmas[i - 1], &a_s1, &s_s1);
      const float lam_s1 = logf(a_s1) - logf(s_s1);
      const float r0 = (lam_s0 - lam_s1) / hh;
      c.inv_r0 = 1.0f / r0;
    }
    if (algo == VV_ALGO_DPMPP) {
      c.cx = s_t / s_s0;
      c.c0 = -(a_t * (expf(-hh) - 1.0f));
      c.c1 = (c.order == 2) ? 0.5f * c.c0 : 0.f;
      c.cn = 0.f;
    } else {
      c.cx = s_t / s_s0 * expf(-hh);
      c.c0 = a_t * (1.0f - expf(-2.0f * hh));
      c.c1 = (c.order == 2) ? 0.5f * c.c0 : 0.f;
      c.cn = s_t * sqrtf(1.0f - expf(-2.0f * hh));
    }
    co[i] = c;
  }
  h->coefs_host = co;
  VV_TRY(h->coefs.alloc(sizeof(StepCoef) * S));
  VV_TRY(h->temb_table.alloc(static_cast<size_t>(S) * h->H * 4));
  DevBuf tv, tfreq, th;
  VV_TRY(tv.alloc(static_cast<size_t>(S) * 4));
  VV_TRY(tfreq.alloc(static_cast<size_t>(S) * h->d.freq_dim * 2));
  VV_TRY(th.alloc(static_cast<size_t>(S) * h->H * 2));
  VV_CHECK_CUDA(cudaMemcpyAsync(h->coefs.p, co.data(), sizeof(StepCoef) * S, cudaMemcpyHostToDevice, st));
  VV_CHECK_CUDA(cudaMemcpyAsync(tv.p, t_values, static_cast<size_t>(S) * 4, cudaMemcpyHostToDevice, st));
  VV_TRY(t_embed_rows(h, tv.as<float>(), S, tfreq.as<bf16>(), th.as<bf16>(), h->temb_table.as<float>(), st));
  VV_COUNT_SYNC();
  VV_CHECK_CUDA(cudaStreamSynchronize(st));   // temporaries are freed on return
  {  // scratch for the all-steps adaLN table (head_modulation), sized here so sample() never allocates
    const size_t rows = static_cast<size_t>(h->d.max_rows) * S;
    if (rows * h->NADA * sizeof(bf16) <= MOD_TABLE_BUDGET) {
      VV_TRY(h->sc.ensure(rows * h->H * sizeof(bf16)));
      VV_TRY(h->mod.ensure(rows * h->NADA * sizeof(bf16)));
    }
  }
  h->S = S;
  h->algo = algo;
  h->chain_S = -1;   // programs embed the schedule: rebuilt on the next sample     // coefficients / tables changed: rebuild the persistent program on next use
  return VV_OK;
}

int vv_head_forward(VVHead* h, const void* z, const float* t, const void* cond, void* out, int M,
                    int io_dtype, void* stream) {
  VV_REQUIRE(h && z && t && cond && out, "vv_head_forward: null argument");
  VV_REQUIRE(M >= 1 && M <= h->d.max_rows, "vv_head_forward: M=%d exceeds max_rows=%d", M, h->d.max_rows);
  VV_REQUIRE(io_dtype == VV_DTYPE_BF16 || io_dtype == VV_DTYPE_F32, "vv_head_forward: io_dtype");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int H = h->H, L = h->L;
  VV_TRY(to_bf16(z, h->zb.as<bf16>(), static_cast<size_t>(M) * L, io_dtype, st));
  VV_TRY(to_bf16(cond, h->condb.as<bf16>(), static_cast<size_t>(M) * H, io_dtype, st));
  VV_TRY(head_gemm(h->condb.as<bf16>(), H, h->w_cond.as<bf16>(), H, M, H, H, GM_STORE_F32, h->cproj.p, H,
                   nullptr, 0, nullptr, 0, nullptr, st));
  VV_TRY(t_embed_rows(h, t, M, h->tfreq.as<bf16>(), h->th.as<bf16>(), h->temb_rows.as<float>(), st));
  int nsv = 1;
  VV_TRY(head_modulation(h, M, 1, h->temb_rows.as<float>(), H, 0, st));
  VV_TRY(head_core(h, M, h->mod.as<bf16>(), &nsv, st));
  const size_t MN = static_cast<size_t>(M) * L;
  k_store_out<<<grid1d(MN, 256), 256, 0, st>>>(h->vpart.as<float>(), nsv, MN, out, io_dtype);
  VV_COUNT_LAUNCH();
  VV_CHECK_CUDA(cudaGetLastError());
  return VV_OK;
}

int vv_head_sample(VVHead* h, const void* cond, const void* neg, const float* noise,
                   const float* sde_noise, float cfg, void* out_latent, int n, int io_dtype,
                   void* stream) {
  VV_REQUIRE(h && cond && neg && noise && out_latent, "vv_head_sample: null argument");
  if (h->S <= 0) {
    vv_set_error("vv_head_sample: call vv_head_set_schedule first");
    return VV_ERR_STATE;
  }
  VV_REQUIRE(n >= 1 && 2 * n <= h->d.max_rows, "vv_head_sample: 2n=%d exceeds max_rows=%d", 2 * n, h->d.max_rows);
  VV_REQUIRE(h->algo != VV_ALGO_SDE_DPMPP || sde_noise != nullptr, "vv_head_sample: sde_noise required for sde-dpmsolver++");
  VV_REQUIRE(io_dtype == VV_DTYPE_BF16 || io_dtype == VV_DTYPE_F32, "vv_head_sample: io_dtype");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int H = h->H, L = h->L, M = 2 * n;
  // condition = cat([cond, neg])  (modeling_vibevoice_inference.py:715)
  VV_TRY(to_bf16(cond, h->condb.as<bf16>(), static_cast<size_t>(n) * H, io_dtype, st));
  VV_TRY(to_bf16(neg, h->condb.as<bf16>() + static_cast<size_t>(n) * H, static_cast<size_t>(n) * H, io_dtype, st));
  // cond_proj(condition) is step-invariant: once per frame (:273)
  VV_TRY(head_gemm(h->condb.as<bf16>(), H, h->w_cond.as<bf16>(), H, M, H, H, GM_STORE_F32, h->cproj.p, H,
                   nullptr, 0, nullptr, 0, nullptr, st));
  const size_t nL = static_cast<size_t>(n) * L;
  // the modulation of all S steps in one GEMM when the table fits the budget, else step by step
  const size_t mod_all_bytes = static_cast<size_t>(M) * h->S * h->NADA * sizeof(bf16);
  const bool hoist = mod_all_bytes <= MOD_TABLE_BUDGET;
  if (hoist) VV_TRY(head_modulation(h, M, h->S, h->temb_table.as<float>(), 0, H, st));
  VV_CHECK_CUDA(vv_launch(k_init_state, grid1d(nL, 256), dim3(256), 0, st, 1, true, noise, h->zstate.as<float>(),
                          h->zb.as<bf16>(), n, L));
  // VibeVoice-Large (H = 3584: 1.28 GB of weights per solver step): in round 1 the per-launch path won by ~4 % (2.78 vs
  // 2.89 ms per frame at n = 1); with the round-2 engine the persistent kernel wins there too (2.36 vs 2.72 ms at n = 1 =
  // 5.4 TB/s, 2.98 vs 3.74 ms at n = 64; scripts/time_large_head.py), so it is the path for every head it can stage
  const int chain_opt = vv_option_head_chain();
  if (hoist && M <= vv_option_head_chain_rows() && chain_opt >= 1) {
    // all S solver steps in one persistent cooperative launch (vv_chain.cu)
    if (h->chain_S != h->S || h->chain_mod != h->mod.p) {
      VV_COUNT_SYNC();
      VV_CHECK_CUDA(cudaStreamSynchronize(st));      // programs in flight may still reference the old tables
      h->drop_chains();
      h->chain_S = h->S; h->chain_mod = h->mod.p;
    }
    auto it = h->chains.find(n);
    if (it == h->chains.end()) {
      if (static_cast<int>(h->chains.size()) >= VVHead::MAX_CHAINS) {
        VV_COUNT_SYNC();
        VV_CHECK_CUDA(cudaStreamSynchronize(st));
        h->drop_chains();
      }
      ChainProgram* cp = chain_create();
      if (head_build_chain(h, n, cp) != VV_OK) {     // e.g. a schedule too long for the program to fit in shared memory:
        chain_destroy(cp);                           // this batch size runs launch by launch instead
        cp = nullptr;
      }
      it = h->chains.emplace(n, cp).first;
    }
    if (it->second != nullptr) {
      ChainCall call{};
      call.cfg = cfg; call.sde_noise = sde_noise; call.out = out_latent; call.out_dtype = io_dtype;
      VV_TRY(chain_launch(it->second, call, st));
      VV_CHECK_CUDA(cudaGetLastError());
      return VV_OK;
    }
  }
  for (int s = 0; s < h->S; ++s) {
    int nsv = 1;
    const bf16* mod = h->mod.as<bf16>();
    if (hoist) mod += static_cast<size_t>(s) * M * h->NADA;
    else VV_TRY(head_modulation(h, M, 1, h->temb_table.as<float>() + static_cast<size_t>(s) * H, 0, 0, st));
    VV_TRY(head_core(h, M, mod, &nsv, st));
    const bool last = (s == h->S - 1);
    VV_CHECK_CUDA(vv_launch(k_cfg_update, grid1d(nL, 256), dim3(256), 0, st, 1, true, h->vpart.as<float>(), nsv,
                            h->zstate.as<float>(), h->prev_x0.as<float>(), h->zb.as<bf16>(),
                            h->coefs.as<StepCoef>(), s, cfg,
                            sde_noise ? sde_noise + static_cast<size_t>(s) * nL : static_cast<const float*>(nullptr),
                            last ? out_latent : static_cast<void*>(nullptr), io_dtype, n, L));
  }
  VV_CHECK_CUDA(cudaGetLastError());
  return VV_OK;
}

}  // extern "C"
