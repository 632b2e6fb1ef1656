// vv_dec.cu — sigma-VAE acoustic tokenizer decoder + in-place streaming conv cache.
//
// Reference: TokenizerDecoder (modular_vibevoice_tokenizer.py:816-951), SConv1d (:258-418),
// SConvTranspose1d (:421-576), Block1D (:620-684, inlined at :924-944), ConvRMSNorm (:77-91),
// FFN (:579-596), VibeVoiceTokenizerStreamingCache (:193-256),
// VibeVoiceAcousticTokenizerModel.decode (:1099-1108).
//
// Layout: activations are channel-last bf16 [n, T, C] (the reference is [n, C, T] and permutes
// around every FFN — :937-939).  Channel-last makes
//   * ConvRMSNorm a row reduction, the depthwise conv a coalesced 7-tap stencil over rows,
//   * the FFN a plain [n*T, C] x [C, 4C] GEMM (no permute copies),
//   * the transposed conv (k = 2*stride) ONE GEMM: A = [x[t] | x[t-1]]  (K = 2*Cin),
//     W' = [stride*Cout, 2*Cin]; its row-major output [n*T, stride*Cout] *is* the channel-last
//     tensor [n, T*stride, Cout] (y[t*s+j, co] = b + x[t]·W[:,co,j] + x[t-1]·W[:,co,j+s]),
//   * the dense stem conv an im2col GEMM with K = 7*vae_dim.
// Streaming cache: ONE arena [2][max_slots + 1][slot_elems] bf16 (the extra slot absorbs out-of-range indices, k_check_slots), updated in place by the layer
// kernels.  Each layer keeps exactly the history it needs (6 rows for k=7 convs, 1 row for a
// transposed conv — the reference keeps k-1 = 15/9/7/3 input frames and recomputes them,
// :447,:539-547).  Two copies + a per-slot parity bit make a step race-free: a layer's first
// tile reads history while its last tile writes the new one.  All-zero history == "no
// history" == set_to_zero (:234-241).  Offline decode is the same code path run on a private
// zeroed cache, chunked in time, so there is a single implementation to validate.
#include <algorithm>
#include <climits>
#include <vector>

#include "../../include/vvb200.h"
#include <map>
#include <tuple>

#include "vv_block.cuh"
#include "vv_chain.cuh"
#include "vv_common.cuh"
#include "vv_gemm.cuh"

namespace {

// ---------------------------------------------------------------------------------------
// Stem: latent de-normalisation (modeling_vibevoice_inference.py:651) + im2col for the causal
// k=7 dense conv (SConv1d, :327-382 streaming / :384-418 non-streaming).
//   A[(b,t), j*V + ci] = z[b, t-6+j, ci]   (history rows from the cache)
// ---------------------------------------------------------------------------------------
template <typename T>
__global__ void k_stem_pack(const T* __restrict__ lat, int time_major, float inv_scale, float bias,
                            bf16* __restrict__ A, CacheView cv, size_t off, int n, int Tn, int V,
                            int Ttot, int toff) {
  griddep_wait();
  griddep_launch();
  const int K = (HIST + 1) * V;
  const size_t main_elems = static_cast<size_t>(n) * Tn * K;
  const size_t cache_elems = static_cast<size_t>(n) * HIST * V;
  const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  if (i >= main_elems + cache_elems) return;
  int b, tsrc, ci;
  bf16* dst;
  if (i < main_elems) {
    const int k = static_cast<int>(i % K);
    const size_t row = i / K;
    b = static_cast<int>(row / Tn);
    const int t = static_cast<int>(row % Tn);
    tsrc = t - HIST + k / V;
    ci = k % V;
    dst = A + i;
  } else {                               // new history = last 6 rows of [old history ; new rows]
    const size_t r = i - main_elems;
    ci = static_cast<int>(r % V);
    const int q = static_cast<int>((r / V) % HIST);
    b = static_cast<int>(r / (static_cast<size_t>(V) * HIST));
    tsrc = Tn - HIST + q;
    dst = cache_wr(cv, b, off) + q * V + ci;
  }
  bf16 v;
  if (tsrc >= 0) {
    // the call covers frames [toff, toff+Tn) of a [n, Ttot] latent tensor
    const size_t src = time_major ? (static_cast<size_t>(b) * Ttot + toff + tsrc) * V + ci
                                  : (static_cast<size_t>(b) * V + ci) * Ttot + toff + tsrc;
    v = __float2bfloat16(static_cast<float>(lat[src]) * inv_scale - bias);
  } else {
    v = cache_rd(cv, b, off)[(HIST + tsrc) * V + ci];
  }
  *dst = v;
}

// ---------------------------------------------------------------------------------------
// Transposed-conv operand: A2[(b,t)] = [ x[b,t,:] | x[b,t-1,:] ], history row from the cache;
// new history = x[b,T-1,:].   (SConvTranspose1d, :478-549 / :551-576)
// ---------------------------------------------------------------------------------------
__global__ void k_convtr_pack(const bf16* __restrict__ x, bf16* __restrict__ A2, CacheView cv,
                              size_t off, int n, int Tn, int C, int valid_t, int Cin) {
  // valid_t / Cin (encoder prefill only): a row is C/Cin time steps of Cin channels; steps >= valid_t read as zero
  // — the right "extra padding" of the non-streaming strided conv (:127-133, :393-401).  INT_MAX: no mask.
  griddep_wait();
  griddep_launch();
  const int vpr = C / 8;                               // uint4 vectors per row
  const size_t total = static_cast<size_t>(n) * Tn * vpr;
  const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  if (i >= total) return;
  const int c8 = static_cast<int>(i % vpr);
  const size_t row = i / vpr;
  const int b = static_cast<int>(row / Tn), t = static_cast<int>(row % Tn);
  uint4 cur = *reinterpret_cast<const uint4*>(x + row * C + c8 * 8);
  uint4 prev;
  if (t > 0) prev = *reinterpret_cast<const uint4*>(x + (row - 1) * C + c8 * 8);
  else prev = *reinterpret_cast<const uint4*>(cache_rd(cv, b, off) + c8 * 8);
  if (valid_t != INT_MAX) {
    const long long step = static_cast<long long>(t) * (C / Cin) + (c8 * 8) / Cin;
    if (step >= valid_t) cur = make_uint4(0u, 0u, 0u, 0u);
    if (t > 0 && step - C / Cin >= valid_t) prev = make_uint4(0u, 0u, 0u, 0u);
  }
  bf16* dst = A2 + row * (2 * static_cast<size_t>(C));
  *reinterpret_cast<uint4*>(dst + c8 * 8) = cur;
  *reinterpret_cast<uint4*>(dst + C + c8 * 8) = prev;
  if (t == Tn - 1) *reinterpret_cast<uint4*>(cache_wr(cv, b, off) + c8 * 8) = cur;
}

// The caller's slot indices are device data (generate() passes a CUDA tensor): an index outside the arena must
// neither read nor write another row's history.  Every cached call first copies them through this kernel: an
// out-of-range index is redirected to the arena's spare slot and counted; the count travels to pinned host memory at
// the end of the call and the next call on the cache reports it (no host sync on the hot path).
__global__ void k_check_slots(const long long* __restrict__ in, long long* __restrict__ out, int n, int max_slots,
                              long long* __restrict__ err) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const long long s = in[i];
  const bool ok = s >= 0 && s < max_slots;
  out[i] = ok ? s : static_cast<long long>(max_slots);
  if (!ok && atomicAdd(reinterpret_cast<unsigned long long*>(err), 1ull) == 0ull) err[1] = s;
}
// par[slot] ^= 1 for the rows stepped this call (after every layer wrote its new history).
__global__ void k_cache_flip(int* __restrict__ parity, const long long* __restrict__ slots, int n) {
  griddep_wait();
  griddep_launch();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) parity[slots[i]] ^= 1;
}
// set_to_zero (:234-241): zero both copies of the listed slots (slots == nullptr: slot i = i).
__global__ void k_cache_zero(bf16* __restrict__ base, const long long* __restrict__ slots, int k,
                             size_t copy_stride, size_t slot_elems) {
  const size_t vecs = slot_elems / 8;
  const size_t total = static_cast<size_t>(k) * 2 * vecs;
  size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (; i < total; i += stride) {
    const size_t v = i % vecs;
    const int copy = static_cast<int>((i / vecs) & 1);
    const size_t j = i / (2 * vecs);
    const long long s = slots ? slots[j] : static_cast<long long>(j);
    reinterpret_cast<uint4*>(base + copy * copy_stride + static_cast<size_t>(s) * slot_elems)[v] =
        make_uint4(0u, 0u, 0u, 0u);
  }
}

// ---------------------------------------------------------------------------------------
// Block1D mixer half, fused (:927-936):
//   xn = ConvRMSNorm(x) ; x' = x + gamma * (dwconv7_causal(xn) + b) ; y = ConvRMSNorm_ffn(x')
// One CTA = one utterance x one tile of TT rows (+6 halo rows recomputed from x, or read from
// the cache when they precede this call).  Each thread owns 8 channels (one 16-byte vector);
// LPR = C/8 lanes cover a row; the normalised rows live in shared memory for the 7-tap stencil.
// ---------------------------------------------------------------------------------------
constexpr int DW_THREADS = 256;

__device__ __forceinline__ void ld8f(const float* p, float (&f)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
// 1-D bulk copy global -> shared (TMA engine, no tensor map); completes on `bar` (complete_tx)
__device__ __forceinline__ void bulk_copy_g2s(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

struct DwParams {
  const bf16* x_in; bf16* x_out; bf16* y;
  const float* norm_w; const float* dw_w;   // dw_w: [7][C] tap-major
  const float* dw_b; const float* gamma; const float* ffn_norm_w;
  CacheView cv; size_t off;
  int Tn, C, TT;
  float eps;
};

// sum over the LPR lanes that share a row. LPR <= 32: sub-warp shuffle; else via shared memory.
__device__ __forceinline__ float row_sum(float v, int lpr, float* red, int tid) {
  if (lpr <= 32) {
    for (int o = lpr >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
  }
  v = warp_sum(v);
  __syncthreads();                       // protect `red` from the previous use
  if ((tid & 31) == 0) red[tid >> 5] = v;
  __syncthreads();
  const int wpr = lpr >> 5;              // warps per row
  const int w0 = ((tid >> 5) / wpr) * wpr;
  float s = 0.f;
  for (int i = 0; i < wpr; ++i) s += red[w0 + i];
  return s;
}

__global__ void __launch_bounds__(DW_THREADS) k_block_dw(const DwParams p) {
  griddep_wait();
  griddep_launch();
  extern __shared__ __align__(16) uint8_t dw_smem[];
  bf16* xn = reinterpret_cast<bf16*>(dw_smem);            // [(TT+6)][C]
  __shared__ float red[DW_THREADS / 32];
  const int C = p.C, Tn = p.Tn, TT = p.TT;
  const int b = blockIdx.y;
  const int t0 = blockIdx.x * TT;
  const int tid = threadIdx.x;
  const int lpr = C >> 3;
  const int lane = tid % lpr;                              // channel group
  const int rslot = tid / lpr;                             // row slot within a pass
  const int rpp = DW_THREADS / lpr;                        // rows per pass (>= 1: C <= 2048)
  const int c0 = lane * 8;
  const bf16* xin = p.x_in + static_cast<size_t>(b) * Tn * C;
  const int rows_here = min(TT, Tn - t0);

  // ---- phase 1: normalised rows (time t0-6 .. t0+rows_here-1) into shared memory ----
  {
    float wn[8];
    {
      const float4 a = __ldg(reinterpret_cast<const float4*>(p.norm_w + c0));
      const float4 c = __ldg(reinterpret_cast<const float4*>(p.norm_w + c0) + 1);
      wn[0] = a.x; wn[1] = a.y; wn[2] = a.z; wn[3] = a.w; wn[4] = c.x; wn[5] = c.y; wn[6] = c.z; wn[7] = c.w;
    }
    const int nrows = rows_here + HIST;
    const int passes = (nrows + rpp - 1) / rpp;
    for (int ps = 0; ps < passes; ++ps) {
      const int r = ps * rpp + rslot;
      const int tg = t0 - HIST + r;
      const bool live = (r < nrows) && tid < rpp * lpr;
      float f[8];
      float ss = 0.f;
      uint4 hv = make_uint4(0u, 0u, 0u, 0u);
      const bool from_x = live && tg >= 0;
      if (from_x) {
        unpack_bf16x8(*reinterpret_cast<const uint4*>(xin + static_cast<size_t>(tg) * C + c0), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) ss += f[i] * f[i];
      } else if (live) {
        hv = *reinterpret_cast<const uint4*>(cache_rd(p.cv, b, p.off) + (HIST + tg) * C + c0);
      }
      ss = row_sum(ss, lpr, red, tid);
      if (from_x) {
        const float rms = rsqrtf(ss / static_cast<float>(C) + p.eps);
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = f[i] * rms * wn[i];
        hv = pack_bf16x8(f);
      }
      if (live) *reinterpret_cast<uint4*>(xn + static_cast<size_t>(r) * C + c0) = hv;
    }
  }
  __syncthreads();

  // ---- new history: normalised rows of time Tn-6 .. Tn-1 (the tile that holds the last row) ----
  if (t0 + TT >= Tn) {
    bf16* dst = cache_wr(p.cv, b, p.off);
    for (int i = tid; i < HIST * lpr; i += DW_THREADS) {
      const int q = i / lpr, l = i % lpr;
      const int r = Tn - t0 + q;                           // smem row of time Tn-6+q
      *reinterpret_cast<uint4*>(dst + q * C + l * 8) =
          *reinterpret_cast<const uint4*>(xn + static_cast<size_t>(r) * C + l * 8);
    }
  }

  // ---- phase 2: 7-tap depthwise conv, layer-scale residual, FFN norm ----
  float wt[HIST + 1][8], bb[8], gm[8], wf[8];
#pragma unroll
  for (int j = 0; j <= HIST; ++j) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p.dw_w + static_cast<size_t>(j) * C + c0));
    const float4 c = __ldg(reinterpret_cast<const float4*>(p.dw_w + static_cast<size_t>(j) * C + c0) + 1);
    wt[j][0] = a.x; wt[j][1] = a.y; wt[j][2] = a.z; wt[j][3] = a.w;
    wt[j][4] = c.x; wt[j][5] = c.y; wt[j][6] = c.z; wt[j][7] = c.w;
  }
  {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p.dw_b + c0)), c = __ldg(reinterpret_cast<const float4*>(p.dw_b + c0) + 1);
    bb[0] = a.x; bb[1] = a.y; bb[2] = a.z; bb[3] = a.w; bb[4] = c.x; bb[5] = c.y; bb[6] = c.z; bb[7] = c.w;
    const float4 g = __ldg(reinterpret_cast<const float4*>(p.gamma + c0)), g2 = __ldg(reinterpret_cast<const float4*>(p.gamma + c0) + 1);
    gm[0] = g.x; gm[1] = g.y; gm[2] = g.z; gm[3] = g.w; gm[4] = g2.x; gm[5] = g2.y; gm[6] = g2.z; gm[7] = g2.w;
    const float4 w = __ldg(reinterpret_cast<const float4*>(p.ffn_norm_w + c0)), w2 = __ldg(reinterpret_cast<const float4*>(p.ffn_norm_w + c0) + 1);
    wf[0] = w.x; wf[1] = w.y; wf[2] = w.z; wf[3] = w.w; wf[4] = w2.x; wf[5] = w2.y; wf[6] = w2.z; wf[7] = w2.w;
  }
  const int passes2 = (rows_here + rpp - 1) / rpp;
  for (int ps = 0; ps < passes2; ++ps) {
    const int r = ps * rpp + rslot;
    const bool live = (r < rows_here) && tid < rpp * lpr;
    const int t = t0 + r;
    float o[8];
    float ss = 0.f;
    if (live) {
      float acc[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = bb[i];
#pragma unroll
      for (int j = 0; j <= HIST; ++j) {
        float v[8];
        unpack_bf16x8(*reinterpret_cast<const uint4*>(xn + static_cast<size_t>(r + j) * C + c0), v);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = fmaf(wt[j][i], v[i], acc[i]);
      }
      float xr[8];
      unpack_bf16x8(*reinterpret_cast<const uint4*>(xin + static_cast<size_t>(t) * C + c0), xr);
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = xr[i] + gm[i] * acc[i];
      const uint4 ov = pack_bf16x8(o);
      *reinterpret_cast<uint4*>(p.x_out + (static_cast<size_t>(b) * Tn + t) * C + c0) = ov;
      unpack_bf16x8(ov, o);                 // the FFN norm sees the stored (bf16) residual stream
#pragma unroll
      for (int i = 0; i < 8; ++i) ss += o[i] * o[i];
    }
    ss = row_sum(ss, lpr, red, tid);
    if (live) {
      const float rms = rsqrtf(ss / static_cast<float>(C) + p.eps);
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = o[i] * rms * wf[i];
      *reinterpret_cast<uint4*>(p.y + (static_cast<size_t>(b) * Tn + t) * C + c0) = pack_bf16x8(o);
    }
  }
}

// Same mixer with the input tile brought into shared memory by bulk copies (cp.async.bulk, one mbarrier) before any
// arithmetic.  k_block_dw reads its rows pass by pass with register loads: (TT + 6) / rows-per-pass dependent L2 round
// trips for the norm, as many again for the residual - 11-15 us per launch for a few MB (vv_set_option("launch_log")).
// Here a CTA pays one transfer latency; both norms, the stencil and the residual then run out of shared memory.
// Arithmetic and its order are k_block_dw's: the two kernels agree bit for bit (tests/test_decoder_gpu.py).
// shared memory: raw x rows [(TT + 6)][C] | normalised rows [(TT + 6)][C]   (rows of time t0 - 6 .. t0 + rows - 1)
__global__ void __launch_bounds__(DW_THREADS) k_block_dw_bulk(const DwParams p) {
  extern __shared__ __align__(16) uint8_t dw_smem[];
  __shared__ float red[DW_THREADS / 32];
  __shared__ __align__(8) uint64_t bar;
  const int C = p.C, Tn = p.Tn, TT = p.TT;
  bf16* xs = reinterpret_cast<bf16*>(dw_smem);
  bf16* xn = xs + static_cast<size_t>(TT + HIST) * C;
  const int b = blockIdx.y;
  const int t0 = blockIdx.x * TT;
  const int tid = threadIdx.x;
  const int lpr = C >> 3;
  const int lane = tid % lpr;
  const int rslot = tid / lpr;
  const int rpp = DW_THREADS / lpr;
  const int c0 = lane * 8;
  const bf16* xin = p.x_in + static_cast<size_t>(b) * Tn * C;
  const int rows_here = min(TT, Tn - t0);
  const int nrows = rows_here + HIST;
  const int n_hist = max(0, HIST - t0);                       // leading rows that precede this call: from the cache, normalised already
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  // per-channel parameters do not depend on the kernel before this one: fetched ahead of the dependency wait (PDL)
  float wn[8], wt[HIST + 1][8], bb[8], gm[8], wf[8];
  ld8f(p.norm_w + c0, wn);
#pragma unroll
  for (int j = 0; j <= HIST; ++j) ld8f(p.dw_w + static_cast<size_t>(j) * C + c0, wt[j]);
  ld8f(p.dw_b + c0, bb);
  ld8f(p.gamma + c0, gm);
  ld8f(p.ffn_norm_w + c0, wf);
  griddep_wait();
  griddep_launch();
  if (tid == 0) {
    const uint32_t hist_bytes = static_cast<uint32_t>(n_hist) * C * 2, x_bytes = static_cast<uint32_t>(nrows - n_hist) * C * 2;
    mbar_arrive_expect_tx(&bar, hist_bytes + x_bytes);
    if (n_hist > 0) bulk_copy_g2s(xn, cache_rd(p.cv, b, p.off) + static_cast<size_t>(HIST - n_hist) * C, hist_bytes, &bar);
    bulk_copy_g2s(xs + static_cast<size_t>(n_hist) * C, xin + static_cast<size_t>(t0 - HIST + n_hist) * C, x_bytes, &bar);
  }
  __syncthreads();                                            // the barrier's initialisation
  mbar_wait(&bar, 0);

  // ---- phase 1: normalise the rows that came from x ----
  {
    const int passes = (nrows - n_hist + rpp - 1) / rpp;
    for (int ps = 0; ps < passes; ++ps) {
      const int r = n_hist + ps * rpp + rslot;
      const bool live = (r < nrows) && tid < rpp * lpr;
      float f[8];
      float ss = 0.f;
      if (live) {
        unpack_bf16x8(*reinterpret_cast<const uint4*>(xs + static_cast<size_t>(r) * C + c0), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) ss += f[i] * f[i];
      }
      ss = row_sum(ss, lpr, red, tid);
      if (live) {
        const float rms = rsqrtf(ss / static_cast<float>(C) + p.eps);
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = f[i] * rms * wn[i];
        *reinterpret_cast<uint4*>(xn + static_cast<size_t>(r) * C + c0) = pack_bf16x8(f);
      }
    }
  }
  __syncthreads();

  // ---- new history: normalised rows of time Tn-6 .. Tn-1 (the tile that holds the last row) ----
  if (t0 + TT >= Tn) {
    bf16* dst = cache_wr(p.cv, b, p.off);
    for (int i = tid; i < HIST * lpr; i += DW_THREADS) {
      const int q = i / lpr, l = i % lpr;
      const int r = Tn - t0 + q;                           // smem row of time Tn-6+q
      *reinterpret_cast<uint4*>(dst + q * C + l * 8) =
          *reinterpret_cast<const uint4*>(xn + static_cast<size_t>(r) * C + l * 8);
    }
  }

  // ---- phase 2: 7-tap depthwise conv, layer-scale residual, FFN norm ----
  const int passes2 = (rows_here + rpp - 1) / rpp;
  for (int ps = 0; ps < passes2; ++ps) {
    const int r = ps * rpp + rslot;
    const bool live = (r < rows_here) && tid < rpp * lpr;
    const int t = t0 + r;
    float o[8];
    float ss = 0.f;
    if (live) {
      float acc[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = bb[i];
#pragma unroll
      for (int j = 0; j <= HIST; ++j) {
        float v[8];
        unpack_bf16x8(*reinterpret_cast<const uint4*>(xn + static_cast<size_t>(r + j) * C + c0), v);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = fmaf(wt[j][i], v[i], acc[i]);
      }
      float xr[8];
      unpack_bf16x8(*reinterpret_cast<const uint4*>(xs + static_cast<size_t>(r + HIST) * C + c0), xr);
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = xr[i] + gm[i] * acc[i];
      const uint4 ov = pack_bf16x8(o);
      *reinterpret_cast<uint4*>(p.x_out + (static_cast<size_t>(b) * Tn + t) * C + c0) = ov;
      unpack_bf16x8(ov, o);                 // the FFN norm sees the stored (bf16) residual stream
#pragma unroll
      for (int i = 0; i < 8; ++i) ss += o[i] * o[i];
    }
    ss = row_sum(ss, lpr, red, tid);
    if (live) {
      const float rms = rsqrtf(ss / static_cast<float>(C) + p.eps);
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = o[i] * rms * wf[i];
      *reinterpret_cast<uint4*>(p.y + (static_cast<size_t>(b) * Tn + t) * C + c0) = pack_bf16x8(o);
    }
  }
}

// ---------------------------------------------------------------------------------------
// Output head: causal dense conv k=7, C -> 1 (:912, :950).  y[t] = b + sum_{j,ci} w[j*C+ci] * x[t-6+j, ci].
// HBM-bound (reads x once, writes one sample per row).  One CTA = one utterance x 256 outputs;
// the (256+6) x C input window is staged in shared memory with an 80-byte row pitch (C = 32)
// so that the 128-bit row reads of consecutive threads hit distinct banks.
// ---------------------------------------------------------------------------------------
constexpr int HC_TT = 256;
__global__ void __launch_bounds__(HC_TT)
k_head_conv(const bf16* __restrict__ x, const float* __restrict__ w, float bias, void* __restrict__ wave,
            int out_dtype, CacheView cv, size_t off, int Tn, int C, size_t out_stride, size_t out_off) {
  griddep_wait();
  griddep_launch();
  extern __shared__ __align__(16) uint8_t hc_smem[];
  const int pitch = C + 8;                                  // elements
  bf16* xs = reinterpret_cast<bf16*>(hc_smem);              // [(HC_TT+6)][pitch]
  float* ws = reinterpret_cast<float*>(hc_smem + static_cast<size_t>(HC_TT + HIST) * pitch * 2);
  const int b = blockIdx.y, t0 = blockIdx.x * HC_TT, tid = threadIdx.x;
  const int rows_here = min(HC_TT, Tn - t0);
  const int vpr = C / 8;
  const bf16* xb = x + static_cast<size_t>(b) * Tn * C;
  for (int i = tid; i < (rows_here + HIST) * vpr; i += HC_TT) {
    const int r = i / vpr, v = i % vpr;
    const int tg = t0 - HIST + r;
    uint4 val;
    if (tg >= 0) val = *reinterpret_cast<const uint4*>(xb + static_cast<size_t>(tg) * C + v * 8);
    else val = *reinterpret_cast<const uint4*>(cache_rd(cv, b, off) + (HIST + tg) * C + v * 8);
    *reinterpret_cast<uint4*>(xs + r * pitch + v * 8) = val;
  }
  for (int i = tid; i < (HIST + 1) * C; i += HC_TT) ws[i] = w[i];
  __syncthreads();
  if (t0 + HC_TT >= Tn) {                                   // new history: rows of time Tn-6 .. Tn-1
    bf16* dst = cache_wr(cv, b, off);
    for (int i = tid; i < HIST * vpr; i += HC_TT) {
      const int q = i / vpr, v = i % vpr;
      *reinterpret_cast<uint4*>(dst + q * C + v * 8) =
          *reinterpret_cast<const uint4*>(xs + (Tn - t0 + q) * pitch + v * 8);
    }
  }
  if (tid < rows_here) {
    float acc = bias;
    for (int j = 0; j <= HIST; ++j) {
      const bf16* row = xs + (tid + j) * pitch;
      const float* wj = ws + j * C;
      for (int v = 0; v < vpr; ++v) {
        float f[8];
        unpack_bf16x8(*reinterpret_cast<const uint4*>(row + v * 8), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc = fmaf(f[i], wj[v * 8 + i], acc);
      }
    }
    const size_t o = static_cast<size_t>(b) * out_stride + out_off + t0 + tid;
    if (out_dtype == VV_DTYPE_BF16) reinterpret_cast<bf16*>(wave)[o] = __float2bfloat16(acc);
    else reinterpret_cast<float*>(wave)[o] = acc;
  }
}

// Split-K tail of the FFN's second Linear at small M (streaming stage 0/1):
//   x_out = x_res + gamma * (sum_z partial_z + bias)      (:940-942)
__global__ void k_reduce_resid(const float* __restrict__ partial, int nsplit, const bf16* __restrict__ xres,
                               const float* __restrict__ gamma, const float* __restrict__ bias,
                               bf16* __restrict__ xout, int M, int C) {
  griddep_wait();
  griddep_launch();
  const int vpr = C / 8;
  const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  if (i >= static_cast<size_t>(M) * vpr) return;
  const int c0 = static_cast<int>(i % vpr) * 8;
  const size_t m = i / vpr;
  float s[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s[k] = bias[c0 + k];
  for (int z = 0; z < nsplit; ++z) {
    const float* p = partial + (static_cast<size_t>(z) * M + m) * C + c0;
    const float4 a = reinterpret_cast<const float4*>(p)[0], c = reinterpret_cast<const float4*>(p)[1];
    s[0] += a.x; s[1] += a.y; s[2] += a.z; s[3] += a.w; s[4] += c.x; s[5] += c.y; s[6] += c.z; s[7] += c.w;
  }
  float xr[8];
  unpack_bf16x8(*reinterpret_cast<const uint4*>(xres + m * C + c0), xr);
#pragma unroll
  for (int k = 0; k < 8; ++k) xr[k] += gamma[c0 + k] * s[k];
  *reinterpret_cast<uint4*>(xout + m * C + c0) = pack_bf16x8(xr);
}

// Split-K tail of a plain Linear / conv-as-GEMM at small M (the encoder's long-K strided convs and head conv):
//   out = sum_z partial_z + bias        (bf16 or fp32 store)
template <typename T>
__global__ void k_reduce_store(const float* __restrict__ partial, int nsplit, const float* __restrict__ bias,
                               T* __restrict__ out, int M, int C) {
  griddep_wait();
  griddep_launch();
  const int vpr = C / 8;
  const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  if (i >= static_cast<size_t>(M) * vpr) return;
  const int c0 = static_cast<int>(i % vpr) * 8;
  const size_t m = i / vpr;
  float s[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s[k] = bias ? bias[c0 + k] : 0.f;
  for (int z = 0; z < nsplit; ++z) {
    const float* p = partial + (static_cast<size_t>(z) * M + m) * C + c0;
    const float4 a = reinterpret_cast<const float4*>(p)[0], c = reinterpret_cast<const float4*>(p)[1];
    s[0] += a.x; s[1] += a.y; s[2] += a.z; s[3] += a.w; s[4] += c.x; s[5] += c.y; s[6] += c.z; s[7] += c.w;
  }
  if constexpr (sizeof(T) == 2) {
    *reinterpret_cast<uint4*>(out + m * C + c0) = pack_bf16x8(s);
  } else {
    float4* o = reinterpret_cast<float4*>(out + m * C + c0);
    o[0] = make_float4(s[0], s[1], s[2], s[3]);
    o[1] = make_float4(s[4], s[5], s[6], s[7]);
  }
}

// ---------------------------------------------------------------------------------------
// Encoder stem (TokenizerEncoder.downsample_layers[0], :731-733): causal k=7 conv 1 -> C0 on raw audio.
//   x[b,t,c] = bias[c] + sum_j w[j][c] * a[b, t-6+j]      history: the last 6 samples (bf16, like the model dtype)
// One thread = one time step x 8 channels.
// ---------------------------------------------------------------------------------------

// Same conv for C = 8 * LPO, LPO in {1, 2, 4, 8} (every shipped configuration: C = 32), without shared memory: LPO lanes
// share an output's channels, each keeps its 8 channels' 7 x 8 weights in registers and walks HCW_R consecutive outputs with
// seven running accumulators - one 16-byte load per 56 FMAs instead of a scalar shared-memory weight load per FMA (the
// staged kernel above ran at ~20 us for 13 MB at n = 64; warm, `vv_set_option("launch_log", 1)`).  Input rows come
// straight from global memory (L2-resident: the last block just wrote them), 64-byte segments per group.
constexpr int HCW_R = 16;                                    // outputs per lane group (+6 warm-up rows)
template <int LPO>
__global__ void __launch_bounds__(256)
k_head_conv_walk(const bf16* __restrict__ x, const float* __restrict__ w, float bias, void* __restrict__ wave,
                 int out_dtype, CacheView cv, size_t off, int Tn, size_t out_stride, size_t out_off) {
  griddep_wait();
  griddep_launch();
  constexpr int C = 8 * LPO;
  constexpr int GROUPS = 256 / LPO;
  const int b = blockIdx.y, tid = threadIdx.x;
  const int l = tid % LPO, g = tid / LPO, c0 = l * 8;
  const int ts = (blockIdx.x * GROUPS + g) * HCW_R;          // first output of this group
  const bf16* xb = x + static_cast<size_t>(b) * Tn * C;
  const bf16* hist = cache_rd(cv, b, off);
  if (blockIdx.x == gridDim.x - 1) {                         // new history: rows of time Tn-6 .. Tn-1
    bf16* dst = cache_wr(cv, b, off);
    for (int i = tid; i < HIST * LPO; i += 256) {
      const int q = i / LPO, v = i % LPO, tg = Tn - HIST + q;
      const bf16* src = (tg >= 0) ? xb + static_cast<size_t>(tg) * C + v * 8 : hist + (HIST + tg) * C + v * 8;
      *reinterpret_cast<uint4*>(dst + q * C + v * 8) = *reinterpret_cast<const uint4*>(src);
    }
  }
  float wr[HIST + 1][8];
#pragma unroll
  for (int j = 0; j <= HIST; ++j) ld8f(w + j * C + c0, wr[j]);
  float acc[HIST + 1];
#pragma unroll
  for (int j = 0; j <= HIST; ++j) acc[j] = 0.f;
  constexpr int NROWS = HCW_R + HIST;
  uint4 rows[NROWS];
#pragma unroll
  for (int k = 0; k < NROWS; ++k) {                          // all loads first: one L2 round trip per thread
    const int tin = ts - HIST + k;
    rows[k] = make_uint4(0u, 0u, 0u, 0u);
    if (ts < Tn) {
      if (tin < 0) rows[k] = *reinterpret_cast<const uint4*>(hist + (HIST + tin) * C + c0);
      else if (tin < Tn) rows[k] = __ldcg(reinterpret_cast<const uint4*>(xb + static_cast<size_t>(tin) * C + c0));
    }
  }
#pragma unroll
  for (int k = 0; k < NROWS; ++k) {
    float f[8];
    unpack_bf16x8(rows[k], f);
#pragma unroll
    for (int j = 0; j <= HIST; ++j) {                        // row k feeds output k - j with tap j
      if (k - j >= 0 && k - j < HCW_R) {
        float d = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) d = fmaf(f[i], wr[j][i], d);
        acc[(k - j) % (HIST + 1)] += d;
      }
    }
    if (k >= HIST) {                                         // output k - 6 is complete
      const int i = k - HIST;
      float v = acc[i % (HIST + 1)];
      acc[i % (HIST + 1)] = 0.f;
#pragma unroll
      for (int o = LPO >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      v += bias;
      if (l == 0 && ts + i < Tn) {
        const size_t oidx = static_cast<size_t>(b) * out_stride + out_off + ts + i;
        if (out_dtype == VV_DTYPE_BF16) reinterpret_cast<bf16*>(wave)[oidx] = __float2bfloat16(v);
        else reinterpret_cast<float*>(wave)[oidx] = v;
      }
    }
  }
}

template <typename T>
__global__ void k_enc_stem(const T* __restrict__ audio, const float* __restrict__ w, const float* __restrict__ bias,
                           bf16* __restrict__ x, CacheView cv, size_t off, int n, int L, int C0, size_t Ltot,
                           size_t loff, int avalid) {
  // avalid: samples of this chunk that exist in the audio buffer (prefill of an arbitrary-length prompt: the rest
  // of the last frame reads as zero and is never dereferenced)
  griddep_wait();
  griddep_launch();
  const int vpr = C0 / 8;
  const size_t total = static_cast<size_t>(n) * L * vpr;
  const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  if (i >= total) return;
  const int c0 = static_cast<int>(i % vpr) * 8;
  const size_t row = i / vpr;
  const int b = static_cast<int>(row / L), t = static_cast<int>(row % L);
  const T* a = audio + static_cast<size_t>(b) * Ltot + loff;
  const bf16* hist = cache_rd(cv, b, off);
  float acc[8];
  ld8f(bias + c0, acc);
#pragma unroll
  for (int j = 0; j <= HIST; ++j) {
    const int ts = t - HIST + j;
    // the reference runs in the model dtype: the chunk it is handed is already bf16 (:673-679)
    const float v = (ts >= 0) ? (ts < avalid ? __bfloat162float(__float2bfloat16(static_cast<float>(a[ts]))) : 0.0f)
                              : __bfloat162float(hist[HIST + ts]);
    float wj[8];
    ld8f(w + static_cast<size_t>(j) * C0 + c0, wj);
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = fmaf(wj[k], v, acc[k]);
  }
  *reinterpret_cast<uint4*>(x + row * C0 + c0) = pack_bf16x8(acc);
  if (t == L - 1 && c0 == 0) {                 // new history = last 6 samples of [old history ; new samples]
    bf16* dst = cache_wr(cv, b, off);
    for (int q = 0; q < HIST; ++q) {
      const int ts = L - HIST + q;
      dst[q] = (ts >= 0) ? __float2bfloat16(ts < avalid ? static_cast<float>(a[ts]) : 0.0f) : hist[HIST + ts];
    }
  }
}

// ------------------------------- weight packing -------------------------------------
// stem: W[co, ci, j] -> Ws[co, j*V + ci]
template <typename T>
__global__ void k_pack_stem_w(const T* __restrict__ w, bf16* __restrict__ out, int Co, int V, int k) {
  const size_t total = static_cast<size_t>(Co) * V * k;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int kk = static_cast<int>(i % (V * k));
    const int co = static_cast<int>(i / (V * k));
    const int j = kk / V, ci = kk % V;
    out[i] = __float2bfloat16(static_cast<float>(w[(static_cast<size_t>(co) * V + ci) * k + j]));
  }
}
// transposed conv: W[ci, co, kk] (kk < 2s) -> Wu[j*Co + co, half*Ci + ci] with kk = j + half*s
template <typename T>
__global__ void k_pack_convtr_w(const T* __restrict__ w, bf16* __restrict__ out, int Ci, int Co, int s) {
  const size_t total = static_cast<size_t>(s) * Co * 2 * Ci;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int col = static_cast<int>(i % (2 * Ci));
    const int row = static_cast<int>(i / (2 * Ci));
    const int half = col / Ci, ci = col % Ci;
    const int j = row / Co, co = row % Co;
    out[i] = __float2bfloat16(static_cast<float>(w[(static_cast<size_t>(ci) * Co + co) * (2 * s) + j + half * s]));
  }
}
// generic strided gather to f32: out[i] = in[(i % inner) * outer_stride + (i / inner)] when transpose,
// used for the depthwise weights [C,1,k] -> [k][C] and the head conv [1,C,k] -> [k*C].
// strided conv (k = 2s): W[co, ci, j] -> Wd[co, (j >= s ? 0 : s*Ci) + (j % s)*Ci + ci]   (current block | previous block)
template <typename T>
__global__ void k_pack_down_w(const T* __restrict__ w, bf16* __restrict__ out, int Co, int Ci, int s) {
  const size_t total = static_cast<size_t>(Co) * Ci * 2 * s;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int j = static_cast<int>(i % (2 * s));
    const int ci = static_cast<int>((i / (2 * s)) % Ci);
    const size_t co = i / (static_cast<size_t>(2) * s * Ci);
    const size_t k = (j >= s ? 0 : static_cast<size_t>(s) * Ci) + static_cast<size_t>(j % s) * Ci + ci;
    out[co * (static_cast<size_t>(2) * s * Ci) + k] = __float2bfloat16(static_cast<float>(w[i]));
  }
}
template <typename T>
__global__ void k_transpose_f32(const T* __restrict__ in, float* __restrict__ out, int rows, int cols) {
  // in: [rows, cols] -> out: [cols, rows]
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * cols) return;
  const int c = i / rows, r = i % rows;
  out[i] = static_cast<float>(in[static_cast<size_t>(r) * cols + c]);
}
template <typename T>
__global__ void k_copy_f32(const T* __restrict__ in, float* __restrict__ out, int n, int repeat) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * repeat) return;
  out[i] = static_cast<float>(in[i % n]);
}
template <typename T>
__global__ void k_copy_bf16(const T* __restrict__ in, bf16* __restrict__ out, size_t n) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x)
    out[i] = __float2bfloat16(static_cast<float>(in[i]));
}

inline dim3 g1(size_t n, int block) { return dim3(static_cast<unsigned>((n + block - 1) / block)); }

}  // namespace

// ------------------------------------------------------------------------------------
// Handles
// ------------------------------------------------------------------------------------
struct BlockW {
  float *norm_w, *dw_w, *dw_b, *gamma, *ffn_norm_w, *b1, *b2, *ffn_gamma;
  bf16 *w1, *w2;
  size_t cache_off;
  bool fused = false;          // C in {32, 64, 128}: one-launch Block1D (vv_block.cu)
  bool ffn_fused = false;      // C == 256: mixer kernel + one-launch FFN (vv_block.cu)
  CUtensorMap tm_w1, tm_w2;    // FFN weights as TMA operands of the fused kernel
};
struct StageW {
  int C = 0, depth = 0, ratio = 1;
  bf16* w_up = nullptr;     // [ratio*C, 2*Cprev]
  float* b_up = nullptr;    // [ratio*C]
  size_t up_cache_off = 0;  // 1 row of Cprev
  std::vector<BlockW> blocks;
};
struct VVCache {
  VVDec* d = nullptr;
  int max_slots = 0;
  int phys_slots = 0;            // max_slots + 1: the last one absorbs rows whose slot index was out of range
  DevBuf arena, parity, iota;
  DevBuf checked;                // [max_slots] the caller's slot indices, clamped into the arena (what the kernels index with)
  DevBuf err;                    // [2] device: number of out-of-range indices seen, the first such index
  long long* err_host = nullptr; // pinned mirror of `err`, refreshed at the end of every call; read by the NEXT call
  ~VVCache() {
    if (err_host) cudaFreeHost(err_host);
  }
};
struct VVDec {
  VVDecDesc desc;
  int V = 0, nst = 0, hop = 1;
  std::vector<StageW> st;
  bf16* w_stem = nullptr; float* b_stem = nullptr; size_t stem_off = 0;
  float* w_head = nullptr; float b_head = 0.f; size_t head_off = 0;
  // encoder handles (vv_enc_create): stem taps [7][C0] fp32, head conv as a GEMM weight [dim, 7*C_last]
  bool encoder = false;
  float* w_stem_f32 = nullptr; bf16* w_head_gemm = nullptr; float* b_head_vec = nullptr;
  size_t slot_elems = 0;
  size_t row_elems = 0;               // max over stages of (upsample_i * C_i): per latent frame
  std::vector<DevBuf*> pool;          // owns every packed weight
  DevBuf xa, xb, ybuf, hid, partial;
  size_t partial_elems = 0;      // fp32 elements of `partial` (dec_partial_elems)
  VVCache* priv = nullptr;            // zero-history cache for offline decode
  // persistent-kernel programs of the wide stages at one new time step per utterance (run_stage_chain), one per
  // (stage, utterances, input buffer)
  std::map<std::tuple<int, int, const void*>, ChainProgram*> stage_chains;
  ~VVDec() {
    for (auto& kv : stage_chains) chain_destroy(kv.second);
    for (DevBuf* b : pool) delete b;
    delete priv;
  }
  template <typename T>
  T* take(size_t n) {
    DevBuf* b = new DevBuf();
    pool.push_back(b);
    if (b->alloc(n * sizeof(T)) != VV_OK) return nullptr;
    return b->as<T>();
  }
};

namespace {

constexpr size_t DEC_PARTIAL_ELEMS = static_cast<size_t>(1) << 24;   // cap of the fp32 split-K scratch (64 MB)
// A handle's scratch: split-K is only planned for the streaming shapes (few rows against long K; every use checks the
// size and falls back to the unsplit kernels), whose partials are at most splits (<= 32) x rows x C (<= 2048) floats with
// rows = max_rows x 8 time steps in the widest split stage: 64 x max_rows x 2048 covers them; at least 4 MB, at most the cap.
static size_t dec_partial_elems(size_t max_rows) {
  return std::min(DEC_PARTIAL_ELEMS, std::max(static_cast<size_t>(1) << 20, static_cast<size_t>(64) * max_rows * 2048));
}
constexpr int DEC_MAX_SPLITS = 16;
constexpr int BLOCK_M_ROWS = 128;      // one GEMM row tile

template <typename T>
int pack_all(VVDec* d, const void* const* w, cudaStream_t st) {
  const VVDecDesc& ds = d->desc;
  const int k = ds.kernel, V = ds.vae_dim;
  int wi = 0;
  size_t off = 0;
  auto f32vec = [&](const void* src, int n, int repeat) -> float* {
    float* dst = d->take<float>(static_cast<size_t>(n) * repeat);
    if (dst) k_copy_f32<T><<<g1(static_cast<size_t>(n) * repeat, 256), 256, 0, st>>>(static_cast<const T*>(src), dst, n, repeat);
    return dst;
  };
  auto bf16mat = [&](const void* src, size_t n) -> bf16* {
    bf16* dst = d->take<bf16>(n);
    if (dst) k_copy_bf16<T><<<1024, 256, 0, st>>>(static_cast<const T*>(src), dst, n);
    return dst;
  };
  const int C0 = ds.channels[0];
  d->w_stem = d->take<bf16>(static_cast<size_t>(C0) * V * k);
  if (!d->w_stem) return VV_ERR_CUDA;
  k_pack_stem_w<T><<<1024, 256, 0, st>>>(static_cast<const T*>(w[wi++]), d->w_stem, C0, V, k);
  d->b_stem = f32vec(w[wi++], C0, 1);
  d->stem_off = off; off += static_cast<size_t>(HIST) * V;
  d->st.resize(d->nst);
  for (int i = 0; i < d->nst; ++i) {
    StageW& s = d->st[i];
    s.C = ds.channels[i]; s.depth = ds.depths[i]; s.ratio = (i > 0) ? ds.ratios[i] : 1;
    if (i > 0) {
      const int Ci = ds.channels[i - 1], r = s.ratio;
      s.w_up = d->take<bf16>(static_cast<size_t>(r) * s.C * 2 * Ci);
      if (!s.w_up) return VV_ERR_CUDA;
      k_pack_convtr_w<T><<<1024, 256, 0, st>>>(static_cast<const T*>(w[wi++]), s.w_up, Ci, s.C, r);
      s.b_up = f32vec(w[wi++], s.C, r);
      s.up_cache_off = off; off += Ci;
    }
    const int C = s.C;
    s.blocks.resize(s.depth);
    for (int j = 0; j < s.depth; ++j) {
      BlockW& b = s.blocks[j];
      b.norm_w = f32vec(w[wi++], C, 1);
      b.dw_w = d->take<float>(static_cast<size_t>(C) * k);
      if (!b.dw_w) return VV_ERR_CUDA;
      k_transpose_f32<T><<<g1(static_cast<size_t>(C) * k, 256), 256, 0, st>>>(static_cast<const T*>(w[wi++]), b.dw_w, C, k);
      b.dw_b = f32vec(w[wi++], C, 1);
      b.gamma = f32vec(w[wi++], C, 1);
      b.ffn_norm_w = f32vec(w[wi++], C, 1);
      b.w1 = bf16mat(w[wi++], static_cast<size_t>(4) * C * C);
      b.b1 = f32vec(w[wi++], 4 * C, 1);
      b.w2 = bf16mat(w[wi++], static_cast<size_t>(4) * C * C);
      b.b2 = f32vec(w[wi++], C, 1);
      b.ffn_gamma = f32vec(w[wi++], C, 1);
      if (!b.norm_w || !b.dw_b || !b.gamma || !b.ffn_norm_w || !b.w1 || !b.b1 || !b.w2 || !b.b2 || !b.ffn_gamma)
        return VV_ERR_CUDA;
      b.cache_off = off; off += static_cast<size_t>(HIST) * C;
      if (block_fused_supported(C, 1)) {
        VV_TRY(block_fused_make_tmaps(b.w1, b.w2, C, &b.tm_w1, &b.tm_w2));
        b.fused = true;
      } else if (ffn_fused_supported(C)) {
        VV_TRY(ffn_fused_make_tmaps(b.w1, b.w2, C, &b.tm_w1, &b.tm_w2));
        b.ffn_fused = true;
      }
    }
  }
  const int Cl = ds.channels[d->nst - 1];
  d->w_head = d->take<float>(static_cast<size_t>(Cl) * k);
  if (!d->w_head) return VV_ERR_CUDA;
  k_transpose_f32<T><<<g1(static_cast<size_t>(Cl) * k, 256), 256, 0, st>>>(static_cast<const T*>(w[wi++]), d->w_head, Cl, k);
  float* bh = f32vec(w[wi++], 1, 1);
  if (!bh) return VV_ERR_CUDA;
  d->head_off = off; off += static_cast<size_t>(HIST) * Cl;
  d->slot_elems = (off + 7) / 8 * 8;
  VV_COUNT_SYNC();
  VV_CHECK_CUDA(cudaStreamSynchronize(st));
  VV_CHECK_CUDA(cudaMemcpy(&d->b_head, bh, sizeof(float), cudaMemcpyDeviceToHost));
  return VV_OK;
}

}  // namespace

namespace {

int dec_gemm(GemmArgs& a, cudaStream_t st) {
  a.b_streaming = (a.M <= 256);
  gemm_plan(a, vv_num_sms());
  return gemm_launch(a, st);
}

// out = A . B^T + bias for a GM_STORE_* launch.  A weight-streaming shape with few output tiles and a long K
// (M <= 128 rows against tens of MB of weights: the encoder's last strided convs and its head conv at streaming
// batch sizes) cannot fill the machine with N-tiles alone: split K over CTAs into fp32 partials and sum them in a
// row-wise kernel, as the FFN's second Linear does.  Otherwise one launch with the epilogue store.
int gemm_bias_splitk(VVDec* d, GemmArgs& a, cudaStream_t st) {
  if (a.M <= 4 * BLOCK_M_ROWS && a.N % 8 == 0 && a.ldo == a.N && vv_option_enc_splitk()) {
    GemmArgs probe = a;
    probe.mode = GM_PARTIAL_F32;
    probe.b_streaming = 1;
    probe.swap = 2;
    gemm_plan(probe, vv_num_sms());
    int sp = probe.splits > 32 ? 32 : probe.splits;
    const int kb = vv_ceil_div(a.K, 64);
    while (sp > 1 && vv_ceil_div(kb, sp) * (sp - 1) >= kb) --sp;
    if (sp >= (a.M <= BLOCK_M_ROWS ? 4 : 2) && static_cast<size_t>(sp) * a.M * a.N <= d->partial_elems) {
      GemmArgs g = a;
      g.mode = GM_PARTIAL_F32; g.out = d->partial.p; g.ldo = a.N; g.bias = nullptr; g.bn = probe.bn; g.splits = sp;
      g.swap = probe.swap;
      VV_TRY(dec_gemm(g, st));
      const unsigned grid = static_cast<unsigned>((static_cast<size_t>(a.M) * a.N / 8 + 255) / 256);
      if (a.mode == GM_STORE_BF16)
        VV_CHECK_CUDA(vv_launch(k_reduce_store<bf16>, dim3(grid), dim3(256), 0, st, 1, true,
                                static_cast<const float*>(d->partial.as<float>()), sp, static_cast<const float*>(a.bias),
                                static_cast<bf16*>(a.out), a.M, a.N));
      else
        VV_CHECK_CUDA(vv_launch(k_reduce_store<float>, dim3(grid), dim3(256), 0, st, 1, true,
                                static_cast<const float*>(d->partial.as<float>()), sp, static_cast<const float*>(a.bias),
                                static_cast<float*>(a.out), a.M, a.N));
      return VV_OK;
    }
  }
  return dec_gemm(a, st);
}

// A wide stage at ONE new time step per utterance (the streaming frame-step: rows = utterances <= 128): its blocks'
// launches are weight streams of 33 MB that each pay ~7 us of fixed cost around ~5 us of streaming
// (DESIGN.md section 7).  Run the whole stage as ONE cooperative launch of the phase engine instead (vv_chain.cu):
// per block  mixer row phase (sums the previous block's split-K partials on the way in) | FFN1 GEMM + GELU |
// FFN2 GEMM, split-K partials  - three grid barriers per block, weights prefetched across them.  The last block's
// partials are reduced by k_reduce_resid after the launch.
int run_stage_chain(VVDec* d, int stage_idx, const StageW& s, const CacheView& cv, int n, bf16*& xa, bf16*& xb, bf16* yb,
                    bf16* hid, cudaStream_t st) {
  const int C = s.C, rows = n;
  const int sms = vv_num_sms();
  int sp = sms / vv_ceil_div(C, 64);
  const int kb = vv_ceil_div(4 * C, 64);
  sp = std::max(1, std::min({sp, kb / 4, DEC_MAX_SPLITS}));
  while (sp > 1 && vv_ceil_div(kb, sp) * (sp - 1) >= kb) --sp;
  VV_REQUIRE(static_cast<size_t>(sp) * rows * C <= d->partial_elems, "stage chain: split-K scratch");
  const auto key = std::make_tuple(stage_idx, n, static_cast<const void*>(xa));
  auto it = d->stage_chains.find(key);
  if (it == d->stage_chains.end()) {
    ChainProgram* prog = chain_create();
    chain_set_kind(prog, 2);
    int rc = VV_OK;
    bf16* res = nullptr;              // buffer holding the previous block's x'
    bf16* in = xa;                    // materialised input of block 0
    bf16* other = xb;
    for (int j = 0; j < s.depth && rc == VV_OK; ++j) {
      const BlockW& b = s.blocks[j];
      ChainDw w{};
      w.norm_w = b.norm_w; w.dw_w = b.dw_w; w.dw_b = b.dw_b; w.gamma = b.gamma; w.ffn_norm_w = b.ffn_norm_w;
      w.off = b.cache_off; w.rows = rows; w.C = C; w.eps = d->desc.eps; w.y = yb;
      bf16* xprime;
      if (j == 0) {
        w.x_in = in; w.x_out = other; xprime = other;
      } else {
        const BlockW& pb = s.blocks[j - 1];
        w.part = d->partial.as<float>(); w.nsplit = sp; w.xres = res; w.gamma2 = pb.ffn_gamma; w.bias2 = pb.b2;
        xprime = (res == xa) ? xb : xa;
        w.x_out = xprime;
      }
      rc = chain_add_dwmix(prog, w);
      res = xprime;
      GemmArgs a1;
      a1.A = yb; a1.lda = C; a1.B = b.w1; a1.ldb = C; a1.M = rows; a1.N = 4 * C; a1.K = C;
      a1.mode = GM_GELU_BF16; a1.out = hid; a1.ldo = 4 * C; a1.bias = b.b1; a1.splits = 1;
      if (rc == VV_OK) rc = chain_add_gemm(prog, a1);
      GemmArgs a2;
      a2.A = hid; a2.lda = 4 * C; a2.B = b.w2; a2.ldb = 4 * C; a2.M = rows; a2.N = C; a2.K = 4 * C;
      a2.mode = GM_PARTIAL_F32; a2.out = d->partial.p; a2.ldo = C; a2.splits = sp;
      if (rc == VV_OK) rc = chain_add_gemm(prog, a2);
    }
    if (rc == VV_OK) {
      VV_COUNT_SYNC();
      VV_CHECK_CUDA(cudaStreamSynchronize(st));
      rc = chain_finalize(prog);
    }
    if (rc != VV_OK) { chain_destroy(prog); return rc; }
    it = d->stage_chains.emplace(key, prog).first;
  }
  ChainCall call{};
  call.cv = cv;
  VV_TRY(chain_launch(it->second, call, st));
  // x' of the last block: block 0 writes xb, block 1 xa, ... ; the reduced output goes to the other buffer
  bf16* res = (s.depth % 2 == 1) ? xb : xa;
  bf16* out = (res == xa) ? xb : xa;
  const BlockW& lb = s.blocks[s.depth - 1];
  VV_CHECK_CUDA(vv_launch(k_reduce_resid, g1(static_cast<size_t>(rows) * C / 8, 256), dim3(256), 0, st, 1, false,
                          static_cast<const float*>(d->partial.as<float>()), sp, static_cast<const bf16*>(res),
                          static_cast<const float*>(lb.ffn_gamma), static_cast<const float*>(lb.b2), out, rows, C));
  if (out != xa) { bf16* t = xa; xa = xb; xb = t; }
  return VV_OK;
}

// A wide stage (C >= 512) at SEVERAL new time steps per utterance (streaming frame-step behind the first up-convolutions:
// C = 1024 at 8 steps, C = 512 at 40 steps per utterance; rows = n * Tc, a few 128-row tiles): each block used to be
// mixer kernel + FFN1 GEMM + FFN2 GEMM (+ split-K reduce), 13-25 us per launch of which most is fixed cost
// (profiles/r02_launches_dec_block_pipe.csv: 47-58 us per block against ~10 us of work).  Run the whole stage as ONE
// cooperative launch of the phase engine: per block  mixer tiles (CH_DWTILE) | FFN1 + GELU | FFN2 with the layer-scale
// residual in its epilogue - three grid barriers per block, weights prefetched across them.  The stage's tensor stays
// in xa (block input and output), x' lives in xb.
int run_stage_chain_tiled(VVDec* d, int stage_idx, const StageW& s, const CacheView& cv, int n, int Tc, bf16* xa, bf16* xb,
                          bf16* yb, bf16* hid, cudaStream_t st) {
  const int C = s.C, rows = n * Tc;
  const auto key = std::make_tuple(stage_idx, n | (Tc << 16), static_cast<const void*>(xa));
  auto it = d->stage_chains.find(key);
  if (it == d->stage_chains.end()) {
    ChainProgram* prog = chain_create();
    chain_set_kind(prog, 3);
    int rc = VV_OK;
    int TT = 8192 / C;
    TT = std::max(1, std::min(TT, 256));
    for (int j = 0; j < s.depth && rc == VV_OK; ++j) {
      const BlockW& b = s.blocks[j];
      ChainDwTile w{};
      w.x_in = xa; w.x_out = xb; w.y = yb;
      w.norm_w = b.norm_w; w.dw_w = b.dw_w; w.dw_b = b.dw_b; w.gamma = b.gamma; w.ffn_norm_w = b.ffn_norm_w;
      w.off = b.cache_off; w.n = n; w.Tn = Tc; w.C = C; w.TT = TT; w.eps = d->desc.eps;
      rc = chain_add_dwtile(prog, w);
      GemmArgs a1;
      a1.A = yb; a1.lda = C; a1.B = b.w1; a1.ldb = C; a1.M = rows; a1.N = 4 * C; a1.K = C;
      a1.mode = GM_GELU_BF16; a1.out = hid; a1.ldo = 4 * C; a1.bias = b.b1; a1.splits = 1;
      if (rc == VV_OK) rc = chain_add_gemm(prog, a1);
      GemmArgs a2;
      a2.A = hid; a2.lda = 4 * C; a2.B = b.w2; a2.ldb = 4 * C; a2.M = rows; a2.N = C; a2.K = 4 * C;
      a2.mode = GM_RESID_GAMMA_BF16; a2.out = xa; a2.ldo = C; a2.bias = b.b2; a2.gamma = b.ffn_gamma;
      a2.resid = xb; a2.ldr = C; a2.splits = 1;
      if (rc == VV_OK) rc = chain_add_gemm(prog, a2);
    }
    if (rc == VV_OK) {
      VV_COUNT_SYNC();
      VV_CHECK_CUDA(cudaStreamSynchronize(st));
      rc = chain_finalize(prog);
    }
    if (rc != VV_OK) { chain_destroy(prog); return rc; }
    it = d->stage_chains.emplace(key, prog).first;
  }
  ChainCall call{};
  call.cv = cv;
  return chain_launch(it->second, call, st);
}

// The Block1D stack of one stage (:924-944 decoder, :785-806 encoder: the same inlined loop) over `rows` = n * Tc
// channel-last rows.  xa holds the stage input and, on return, its output (xa/xb are swapped as needed).
int run_stage_blocks(VVDec* d, const StageW& s, const CacheView& cv, int n, int Tc, int rows, bf16*& xa, bf16*& xb,
                     bf16* yb, bf16* hid, cudaStream_t st) {
  const int C = s.C;
  if (vv_option_stage_chain() && Tc == 1 && rows <= BLOCK_M_ROWS && rows >= vv_option_stage_chain() && C >= 512 &&
      C <= 2048 && s.depth >= 1 && !((s.blocks[0].fused || s.blocks[0].ffn_fused) && vv_option_fused_block()))
    return run_stage_chain(d, static_cast<int>(&s - d->st.data()), s, cv, n, xa, xb, yb, hid, st);
  if (vv_option_stage_chain() && vv_option_stage_tiled() && Tc > 1 && Tc < 65536 && rows > 0 && rows <= vv_option_stage_tiled() &&
      n >= vv_option_stage_chain() && C >= 512 && C <= 2048 && s.depth >= 1 &&
      !((s.blocks[0].fused || s.blocks[0].ffn_fused) && vv_option_fused_block()))
    return run_stage_chain_tiled(d, static_cast<int>(&s - d->st.data()), s, cv, n, Tc, xa, xb, yb, hid, st);
  {
    int TT = vv_option_dw_rows() > 0 ? vv_option_dw_rows() : 8192 / C;
    if (TT < 1) TT = 1;
    if (TT > 256) TT = 256;
    const size_t dw_smem = static_cast<size_t>(TT + HIST) * C * sizeof(bf16);
    for (int j = 0; j < s.depth; ++j) {
      const BlockW& b = s.blocks[j];
      if (b.fused && vv_option_fused_block()) {
        BlockFusedArgs fa;
        fa.x_in = xa; fa.x_out = xb;
        fa.norm_w = b.norm_w; fa.dw_w = b.dw_w; fa.dw_b = b.dw_b; fa.gamma = b.gamma; fa.ffn_norm_w = b.ffn_norm_w;
        fa.b1 = b.b1; fa.b2 = b.b2; fa.ffn_gamma = b.ffn_gamma;
        fa.tm_w1 = &b.tm_w1; fa.tm_w2 = &b.tm_w2;
        fa.cv = cv; fa.cache_off = b.cache_off;
        fa.n = n; fa.Tn = Tc; fa.C = C; fa.eps = d->desc.eps;
        VV_TRY(block_fused_launch(fa, st));
        bf16* t = xa; xa = xb; xb = t;
        continue;
      }
      DwParams p;
      p.x_in = xa; p.x_out = xb; p.y = yb;
      p.norm_w = b.norm_w; p.dw_w = b.dw_w; p.dw_b = b.dw_b; p.gamma = b.gamma; p.ffn_norm_w = b.ffn_norm_w;
      p.cv = cv; p.off = b.cache_off; p.Tn = Tc; p.C = C; p.TT = TT; p.eps = d->desc.eps;
      if (vv_option_dw_bulk() && 2 * dw_smem <= 96 * 1024)
        VV_CHECK_CUDA(vv_launch(k_block_dw_bulk, dim3(vv_ceil_div(Tc, TT), n), dim3(DW_THREADS), 2 * dw_smem, st, 1, true, p));
      else
        VV_CHECK_CUDA(vv_launch(k_block_dw, dim3(vv_ceil_div(Tc, TT), n), dim3(DW_THREADS), dw_smem, st, 1, true, p));
      // ---- FFN: Linear(C,4C) -> GELU -> Linear(4C,C), layer-scale residual (:936-942) ----
      if (b.ffn_fused && vv_option_fused_block()) {
        FfnFusedArgs fa;
        fa.y = yb; fa.xres = xb; fa.x_out = xa;
        fa.b1 = b.b1; fa.b2 = b.b2; fa.ffn_gamma = b.ffn_gamma;
        fa.tm_w1 = &b.tm_w1; fa.tm_w2 = &b.tm_w2;
        fa.rows = rows; fa.C = C;
        VV_TRY(ffn_fused_launch(fa, st));
        continue;
      }
      GemmArgs a1;
      a1.A = yb; a1.lda = C; a1.B = b.w1; a1.ldb = C; a1.M = rows; a1.N = 4 * C; a1.K = C;
      a1.mode = GM_GELU_BF16; a1.out = hid; a1.ldo = 4 * C; a1.bias = b.b1;
      VV_TRY(dec_gemm(a1, st));
      GemmArgs a2;
      a2.A = hid; a2.lda = 4 * C; a2.B = b.w2; a2.ldb = 4 * C; a2.M = rows; a2.N = C; a2.K = 4 * C;
      // few output tiles + long K (streaming stage 0/1): split K, reduce in a row-wise kernel
      GemmArgs probe = a2;
      probe.mode = GM_PARTIAL_F32;
      probe.swap = 2;
      gemm_plan(probe, vv_num_sms());
      int sp = probe.splits > DEC_MAX_SPLITS ? DEC_MAX_SPLITS : probe.splits;
      if (sp > 1 && static_cast<size_t>(sp) * rows * C <= d->partial_elems) {
        const int kb = vv_ceil_div(a2.K, 64);
        while (sp > 1 && vv_ceil_div(kb, sp) * (sp - 1) >= kb) --sp;
        a2.mode = GM_PARTIAL_F32; a2.out = d->partial.p; a2.ldo = C; a2.bn = probe.bn; a2.splits = sp;
        a2.swap = probe.swap;
        VV_TRY(dec_gemm(a2, st));
        VV_CHECK_CUDA(vv_launch(k_reduce_resid, g1(static_cast<size_t>(rows) * C / 8, 256), dim3(256), 0, st, 1, true,
                                static_cast<const float*>(d->partial.as<float>()), sp, static_cast<const bf16*>(xb),
                                static_cast<const float*>(b.ffn_gamma), static_cast<const float*>(b.b2), xa, rows, C));
      } else {
        a2.mode = GM_RESID_GAMMA_BF16; a2.out = xa; a2.ldo = C; a2.bias = b.b2; a2.gamma = b.ffn_gamma;
        a2.resid = xb; a2.ldr = C;
        VV_TRY(dec_gemm(a2, st));
      }
    }
  }
  return VV_OK;
}

// One pass of the decoder over frames [toff, toff+Tn) of n utterances with cache `c`
// (rows -> slots[]).  TokenizerDecoder.forward (:948-951) / forward_features (:914-946).
int dec_run(VVDec* d, VVCache* c, const long long* slots, const void* lat, int n, int Tn, int Ttot,
            int toff, int time_major, float scale, float bias, void* wave, int io_dtype,
            cudaStream_t st) {
  const int V = d->V;
  CacheView cv;
  cv.base = c->arena.as<bf16>();
  cv.parity = c->parity.as<int>();
  cv.slots = slots;
  cv.copy_stride = static_cast<size_t>(c->phys_slots) * d->slot_elems;
  cv.slot_elems = d->slot_elems;
  bf16* xa = d->xa.as<bf16>();
  bf16* xb = d->xb.as<bf16>();
  bf16* yb = d->ybuf.as<bf16>();
  bf16* hid = d->hid.as<bf16>();
  int rows = n * Tn;     // current number of time rows over the batch
  int Tc = Tn;           // per-utterance length at the current stage
  // ---- stem: de-normalise + im2col, then the dense conv as a GEMM (K = 7*V) ----
  {
    const int K = (HIST + 1) * V;
    const size_t total = static_cast<size_t>(rows) * K + static_cast<size_t>(n) * HIST * V;
    if (io_dtype == VV_DTYPE_BF16)
      VV_CHECK_CUDA(vv_launch(k_stem_pack<bf16>, g1(total, 256), dim3(256), 0, st, 1, true,
                              static_cast<const bf16*>(lat), time_major, 1.0f / scale, bias, hid, cv, d->stem_off, n,
                              Tn, V, Ttot, toff));
    else
      VV_CHECK_CUDA(vv_launch(k_stem_pack<float>, g1(total, 256), dim3(256), 0, st, 1, true,
                              static_cast<const float*>(lat), time_major, 1.0f / scale, bias, hid, cv, d->stem_off, n,
                              Tn, V, Ttot, toff));
    GemmArgs a;
    a.A = hid; a.lda = K; a.B = d->w_stem; a.ldb = K; a.M = rows; a.N = d->st[0].C; a.K = K;
    a.mode = GM_STORE_BF16; a.out = xa; a.ldo = a.N; a.bias = d->b_stem;
    VV_TRY(dec_gemm(a, st));
  }
  for (int i = 0; i < d->nst; ++i) {
    const StageW& s = d->st[i];
    const int C = s.C;
    if (i > 0) {
      // ---- SConvTranspose1d as one GEMM over [x[t] | x[t-1]] ----
      const int Ci = d->st[i - 1].C, r = s.ratio;
      VV_CHECK_CUDA(vv_launch(k_convtr_pack, g1(static_cast<size_t>(rows) * Ci / 8, 256), dim3(256), 0, st, 1, true,
                              static_cast<const bf16*>(xa), hid, cv, s.up_cache_off, n, Tc, Ci, INT_MAX, Ci));
      GemmArgs a;
      a.A = hid; a.lda = 2 * Ci; a.B = s.w_up; a.ldb = 2 * Ci; a.M = rows; a.N = r * C; a.K = 2 * Ci;
      a.mode = GM_STORE_BF16; a.out = xb; a.ldo = a.N; a.bias = s.b_up;
      VV_TRY(dec_gemm(a, st));
      bf16* t = xa; xa = xb; xb = t;
      rows *= r; Tc *= r;
    }
    VV_TRY(run_stage_blocks(d, s, cv, n, Tc, rows, xa, xb, yb, hid, st));
  }
  // ---- head conv (norm is Identity: disable_last_norm, :908-911) ----
  {
    const int C = d->st[d->nst - 1].C;
    const size_t smem = static_cast<size_t>(HC_TT + HIST) * (C + 8) * 2 + static_cast<size_t>(HIST + 1) * C * 4;
    const size_t ostride = static_cast<size_t>(Ttot) * d->hop, ooff = static_cast<size_t>(toff) * d->hop;
    const bf16* xin = xa;
    const float* wh = d->w_head;
    auto walk = [&](auto kern, int lpo) {
      const unsigned gx = static_cast<unsigned>(vv_ceil_div(Tc, (256 / lpo) * HCW_R));
      return vv_launch(kern, dim3(gx, n), dim3(256), 0, st, 1, true, xin, wh, d->b_head, wave, io_dtype, cv, d->head_off, Tc,
                       ostride, ooff);
    };
    if (!vv_option_head_conv_walk() || (C != 8 && C != 16 && C != 32 && C != 64)) {
      VV_CHECK_CUDA(vv_launch(k_head_conv, dim3(vv_ceil_div(Tc, HC_TT), n), dim3(HC_TT), smem, st, 1, true, xin, wh, d->b_head,
                              wave, io_dtype, cv, d->head_off, Tc, C, ostride, ooff));
    } else if (C == 32) {
      VV_CHECK_CUDA(walk(k_head_conv_walk<4>, 4));
    } else if (C == 64) {
      VV_CHECK_CUDA(walk(k_head_conv_walk<8>, 8));
    } else if (C == 16) {
      VV_CHECK_CUDA(walk(k_head_conv_walk<2>, 2));
    } else {
      VV_CHECK_CUDA(walk(k_head_conv_walk<1>, 1));
    }
  }
  VV_CHECK_CUDA(vv_launch(k_cache_flip, g1(n, 128), dim3(128), 0, st, 1, true, c->parity.as<int>(), slots, n));
  VV_CHECK_CUDA(cudaGetLastError());
  return VV_OK;
}

int cache_alloc(VVDec* d, int max_slots, VVCache** out) {
  VVCache* c = new VVCache();
  c->d = d; c->max_slots = max_slots; c->phys_slots = max_slots + 1;
  int rc = c->arena.alloc(static_cast<size_t>(2) * c->phys_slots * d->slot_elems * sizeof(bf16));
  if (rc == VV_OK) rc = c->parity.alloc(static_cast<size_t>(c->phys_slots) * sizeof(int));
  if (rc == VV_OK) rc = c->iota.alloc(static_cast<size_t>(max_slots) * sizeof(long long));
  if (rc == VV_OK) rc = c->checked.alloc(static_cast<size_t>(max_slots) * sizeof(long long));
  if (rc == VV_OK) rc = c->err.alloc(2 * sizeof(long long));
  if (rc == VV_OK && cudaHostAlloc(reinterpret_cast<void**>(&c->err_host), 2 * sizeof(long long), cudaHostAllocDefault) != cudaSuccess) {
    vv_set_error("cache init: cudaHostAlloc failed");
    rc = VV_ERR_CUDA;
  }
  if (rc != VV_OK) { delete c; return rc; }
  c->err_host[0] = 0; c->err_host[1] = 0;
  std::vector<long long> io(max_slots);
  for (int i = 0; i < max_slots; ++i) io[i] = i;
  cudaError_t e = cudaMemset(c->arena.p, 0, c->arena.bytes);
  if (e == cudaSuccess) e = cudaMemset(c->parity.p, 0, c->parity.bytes);
  if (e == cudaSuccess) e = cudaMemset(c->err.p, 0, c->err.bytes);
  if (e == cudaSuccess) e = cudaMemcpy(c->iota.p, io.data(), io.size() * sizeof(long long), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    vv_set_error("cache init: %s", cudaGetErrorString(e));
    delete c;
    return VV_ERR_CUDA;
  }
  *out = c;
  return VV_OK;
}

// Reports what an EARLIER call on this cache found (see k_check_slots), then clamps this call's indices.
int cache_check_slots(VVCache* c, const long long* in, int n, cudaStream_t st, const long long** out, const char* who) {
  if (c->err_host[0] > 0) {
    const long long cnt = c->err_host[0], first = c->err_host[1];
    c->err_host[0] = 0; c->err_host[1] = 0;
    cudaMemsetAsync(c->err.p, 0, c->err.bytes, st);
    vv_set_error("%s: an earlier call on this cache passed %lld slot index(es) outside [0, %d) (first: %lld); those rows used "
                 "a spare slot and their audio is invalid - size the cache for the whole batch (max_slots)",
                 who, cnt, c->max_slots, first);
    return VV_ERR_INVALID;
  }
  k_check_slots<<<(n + 255) / 256, 256, 0, st>>>(in, c->checked.as<long long>(), n, c->max_slots, c->err.as<long long>());
  VV_COUNT_LAUNCH();
  VV_CHECK_CUDA(cudaGetLastError());
  *out = c->checked.as<long long>();
  return VV_OK;
}
int cache_publish_errors(VVCache* c, cudaStream_t st) {
  VV_CHECK_CUDA(cudaMemcpyAsync(c->err_host, c->err.p, 2 * sizeof(long long), cudaMemcpyDeviceToHost, st));
  return VV_OK;
}

}  // namespace

// ------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------
extern "C" {

int vv_dec_create(const VVDecDesc* desc, const void* const* weights, int n_weights, VVDec** out) {
  VV_REQUIRE(desc && weights && out, "vv_dec_create: null argument");
  const int nst = desc->n_stages;
  VV_REQUIRE(nst >= 1 && nst <= VV_DEC_MAX_STAGES, "vv_dec_create: n_stages=%d", nst);
  VV_REQUIRE(desc->kernel == HIST + 1, "vv_dec_create: kernel=%d (only 7 is supported)", desc->kernel);
  VV_REQUIRE(desc->vae_dim > 0 && desc->vae_dim % 8 == 0, "vv_dec_create: vae_dim=%d must be a multiple of 8", desc->vae_dim);
  VV_REQUIRE(desc->max_rows >= 1, "vv_dec_create: max_rows=%d", desc->max_rows);
  int expect = 2 + 2;
  size_t row_elems = 0;
  int up = 1;
  for (int i = 0; i < nst; ++i) {
    const int C = desc->channels[i];
    VV_REQUIRE(C >= 8 && C <= 2048 && (C & (C - 1)) == 0, "vv_dec_create: channels[%d]=%d must be a power of two in [8,2048]", i, C);
    VV_REQUIRE(desc->depths[i] >= 0 && desc->depths[i] <= 64, "vv_dec_create: depths[%d]=%d", i, desc->depths[i]);
    if (i > 0) {
      VV_REQUIRE(desc->ratios[i] >= 1 && desc->ratios[i] <= 64, "vv_dec_create: ratios[%d]=%d", i, desc->ratios[i]);
      up *= desc->ratios[i];
      expect += 2;
    }
    expect += 10 * desc->depths[i];
    const size_t e = static_cast<size_t>(up) * C;
    if (e > row_elems) row_elems = e;
  }
  VV_REQUIRE(n_weights == expect, "vv_dec_create: expected %d weights, got %d", expect, n_weights);
  for (int i = 0; i < n_weights; ++i) VV_REQUIRE(weights[i] != nullptr, "vv_dec_create: weight %d is null", i);
  VV_REQUIRE(desc->weight_dtype == VV_DTYPE_BF16 || desc->weight_dtype == VV_DTYPE_F32, "vv_dec_create: weight_dtype");
  VVDec* d = new VVDec();
  d->desc = *desc; d->V = desc->vae_dim; d->nst = nst; d->hop = up; d->row_elems = row_elems;
  int rc = (desc->weight_dtype == VV_DTYPE_F32) ? pack_all<float>(d, weights, 0) : pack_all<bf16>(d, weights, 0);
  const size_t R = desc->max_rows;
  // hid holds the 4C-wide FFN hidden, the 2*Cin transposed-conv operand and the stem im2col
  size_t hid_elems = 4 * row_elems;
  const size_t stem_elems = static_cast<size_t>(HIST + 1) * d->V;
  if (stem_elems > hid_elems) hid_elems = stem_elems;
  if (rc == VV_OK) rc = d->xa.alloc(R * row_elems * 2);
  if (rc == VV_OK) rc = d->xb.alloc(R * row_elems * 2);
  if (rc == VV_OK) rc = d->ybuf.alloc(R * row_elems * 2);
  if (rc == VV_OK) rc = d->hid.alloc(R * hid_elems * 2);
  d->partial_elems = dec_partial_elems(R);
  if (rc == VV_OK) rc = d->partial.alloc(d->partial_elems * sizeof(float));
  if (rc == VV_OK) {
    cudaError_t e = cudaFuncSetAttribute(k_block_dw, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_block_dw_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) {
      vv_set_error("vv_dec_create: %s", cudaGetErrorString(e));
      rc = VV_ERR_CUDA;
    }
  }
  if (rc != VV_OK) { delete d; return rc; }
  *out = d;
  return VV_OK;
}

void vv_dec_destroy(VVDec* d) { delete d; }

int vv_cache_create(VVDec* d, int max_slots, VVCache** out) {
  VV_REQUIRE(d && out, "vv_cache_create: null argument");
  VV_REQUIRE(max_slots >= 1, "vv_cache_create: max_slots=%d", max_slots);
  return cache_alloc(d, max_slots, out);
}
void vv_cache_destroy(VVCache* c) { delete c; }

int vv_cache_zero(VVCache* c, const int64_t* slot_idx, int k, void* stream) {
  VV_REQUIRE(c && (slot_idx || k == 0), "vv_cache_zero: null argument");
  VV_REQUIRE(k >= 0 && k <= c->max_slots, "vv_cache_zero: k=%d exceeds max_slots=%d", k, c->max_slots);
  if (k == 0) return VV_OK;
  const long long* slots = nullptr;
  VV_TRY(cache_check_slots(c, reinterpret_cast<const long long*>(slot_idx), k, static_cast<cudaStream_t>(stream), &slots,
                           "vv_cache_zero"));
  k_cache_zero<<<256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      c->arena.as<bf16>(), slots, k, static_cast<size_t>(c->phys_slots) * c->d->slot_elems, c->d->slot_elems);
  VV_COUNT_LAUNCH();
  VV_CHECK_CUDA(cudaGetLastError());
  return cache_publish_errors(c, static_cast<cudaStream_t>(stream));
}
int vv_cache_zero_all(VVCache* c, void* stream) {
  VV_REQUIRE(c, "vv_cache_zero_all: null argument");
  VV_CHECK_CUDA(cudaMemsetAsync(c->arena.p, 0, c->arena.bytes, static_cast<cudaStream_t>(stream)));
  return VV_OK;
}

int vv_dec_step(VVDec* d, VVCache* c, const void* latents, const int64_t* slot_idx, int n, int T,
                int time_major, float scale, float bias, void* wave, int io_dtype, void* stream) {
  VV_REQUIRE(d && c && latents && slot_idx && wave, "vv_dec_step: null argument");
  VV_REQUIRE(c->d == d, "vv_dec_step: cache belongs to another decoder");
  VV_REQUIRE(!d->encoder, "vv_dec_step: handle is an encoder");
  VV_REQUIRE(n >= 1 && T >= 1, "vv_dec_step: n=%d T=%d", n, T);
  VV_REQUIRE(n <= c->max_slots, "vv_dec_step: n=%d exceeds cache slots %d", n, c->max_slots);
  VV_REQUIRE(static_cast<long long>(n) * T <= d->desc.max_rows, "vv_dec_step: n*T=%lld exceeds max_rows=%d",
             static_cast<long long>(n) * T, d->desc.max_rows);
  VV_REQUIRE(scale != 0.f, "vv_dec_step: scale must be non-zero");
  VV_REQUIRE(io_dtype == VV_DTYPE_BF16 || io_dtype == VV_DTYPE_F32, "vv_dec_step: io_dtype");
  const long long* slots = nullptr;
  VV_TRY(cache_check_slots(c, reinterpret_cast<const long long*>(slot_idx), n, static_cast<cudaStream_t>(stream), &slots,
                           "vv_dec_step"));
  VV_TRY(dec_run(d, c, slots, latents, n, T, T, 0, time_major, scale, bias, wave, io_dtype, static_cast<cudaStream_t>(stream)));
  return cache_publish_errors(c, static_cast<cudaStream_t>(stream));
}

int vv_dec_decode(VVDec* d, const void* latents, int n, int T, int time_major, float scale, float bias,
                  void* wave, int io_dtype, void* stream) {
  VV_REQUIRE(d && latents && wave, "vv_dec_decode: null argument");
  VV_REQUIRE(n >= 1 && T >= 1, "vv_dec_decode: n=%d T=%d", n, T);
  VV_REQUIRE(n <= d->desc.max_rows, "vv_dec_decode: n=%d exceeds max_rows=%d", n, d->desc.max_rows);
  VV_REQUIRE(!d->encoder, "vv_dec_decode: handle is an encoder");
  VV_REQUIRE(scale != 0.f, "vv_dec_decode: scale must be non-zero");
  VV_REQUIRE(io_dtype == VV_DTYPE_BF16 || io_dtype == VV_DTYPE_F32, "vv_dec_decode: io_dtype");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (d->priv == nullptr || d->priv->max_slots < n) {
    VV_COUNT_SYNC();
    VV_CHECK_CUDA(cudaStreamSynchronize(st));
    delete d->priv;
    d->priv = nullptr;
    VV_TRY(cache_alloc(d, n, &d->priv));
  }
  // zero history (pad_mode='constant' left padding, :398-401), then chunk in time: the cache
  // carries the causal halo from chunk to chunk, so memory stays bounded by max_rows.
  k_cache_zero<<<256, 256, 0, st>>>(d->priv->arena.as<bf16>(), nullptr, n,
                                    static_cast<size_t>(d->priv->phys_slots) * d->slot_elems, d->slot_elems);
  VV_COUNT_LAUNCH();
  const int Tc = d->desc.max_rows / n;
  for (int t0 = 0; t0 < T; t0 += Tc) {
    const int Tn = (T - t0 < Tc) ? (T - t0) : Tc;
    VV_TRY(dec_run(d, d->priv, d->priv->iota.as<long long>(), latents, n, Tn, T, t0, time_major, scale, bias,
                   wave, io_dtype, st));
  }
  return VV_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------
// Semantic-tokenizer encoder: TokenizerEncoder (:687-813) behind
// VibeVoiceSemanticTokenizerModel.encode (:1167-1170), called once per frame on the chunk the decoder
// just produced (modeling_vibevoice_inference.py:673-679).  Mirror image of the decoder: the same
// Block1D stack, the same cache arena; a strided conv with k = 2*stride is ONE GEMM over
// [block[t] | block[t-1]] where block[t] = the `stride` consecutive channel-last rows it consumes
// (one contiguous row of stride*Cin elements), so k_convtr_pack and its 1-row history serve it as is.
// ------------------------------------------------------------------------------------
namespace {

template <typename T>
int pack_all_enc(VVDec* d, const void* const* w, cudaStream_t st) {
  const VVDecDesc& ds = d->desc;
  const int k = ds.kernel;
  int wi = 0;
  size_t off = 0;
  auto f32vec = [&](const void* src, int n) -> float* {
    float* dst = d->take<float>(static_cast<size_t>(n));
    if (dst) k_copy_f32<T><<<g1(static_cast<size_t>(n), 256), 256, 0, st>>>(static_cast<const T*>(src), dst, n, 1);
    return dst;
  };
  auto bf16mat = [&](const void* src, size_t n) -> bf16* {
    bf16* dst = d->take<bf16>(n);
    if (dst) k_copy_bf16<T><<<1024, 256, 0, st>>>(static_cast<const T*>(src), dst, n);
    return dst;
  };
  const int C0 = ds.channels[0];
  d->w_stem_f32 = d->take<float>(static_cast<size_t>(C0) * k);          // [C0, 1, 7] -> [7][C0]
  if (!d->w_stem_f32) return VV_ERR_CUDA;
  k_transpose_f32<T><<<g1(static_cast<size_t>(C0) * k, 256), 256, 0, st>>>(static_cast<const T*>(w[wi++]), d->w_stem_f32, C0, k);
  d->b_stem = f32vec(w[wi++], C0);
  d->stem_off = off; off += 8;                                            // 6 raw samples (padded to 16 bytes)
  d->st.resize(d->nst);
  for (int i = 0; i < d->nst; ++i) {
    StageW& s = d->st[i];
    s.C = ds.channels[i]; s.depth = ds.depths[i]; s.ratio = (i > 0) ? ds.ratios[i] : 1;
    if (i > 0) {
      const int Ci = ds.channels[i - 1], r = s.ratio;
      s.w_up = d->take<bf16>(static_cast<size_t>(s.C) * 2 * r * Ci);
      if (!s.w_up) return VV_ERR_CUDA;
      k_pack_down_w<T><<<1024, 256, 0, st>>>(static_cast<const T*>(w[wi++]), s.w_up, s.C, Ci, r);
      s.b_up = f32vec(w[wi++], s.C);
      s.up_cache_off = off; off += static_cast<size_t>(r) * Ci;          // the previous block of `stride` rows
    }
    const int C = s.C;
    s.blocks.resize(s.depth);
    for (int j = 0; j < s.depth; ++j) {
      BlockW& b = s.blocks[j];
      b.norm_w = f32vec(w[wi++], C);
      b.dw_w = d->take<float>(static_cast<size_t>(C) * k);
      if (!b.dw_w) return VV_ERR_CUDA;
      k_transpose_f32<T><<<g1(static_cast<size_t>(C) * k, 256), 256, 0, st>>>(static_cast<const T*>(w[wi++]), b.dw_w, C, k);
      b.dw_b = f32vec(w[wi++], C);
      b.gamma = f32vec(w[wi++], C);
      b.ffn_norm_w = f32vec(w[wi++], C);
      b.w1 = bf16mat(w[wi++], static_cast<size_t>(4) * C * C);
      b.b1 = f32vec(w[wi++], 4 * C);
      b.w2 = bf16mat(w[wi++], static_cast<size_t>(4) * C * C);
      b.b2 = f32vec(w[wi++], C);
      b.ffn_gamma = f32vec(w[wi++], C);
      if (!b.norm_w || !b.dw_b || !b.gamma || !b.ffn_norm_w || !b.w1 || !b.b1 || !b.w2 || !b.b2 || !b.ffn_gamma)
        return VV_ERR_CUDA;
      b.cache_off = off; off += static_cast<size_t>(HIST) * C;
      if (block_fused_supported(C, 1)) {
        VV_TRY(block_fused_make_tmaps(b.w1, b.w2, C, &b.tm_w1, &b.tm_w2));
        b.fused = true;
      } else if (ffn_fused_supported(C)) {
        VV_TRY(ffn_fused_make_tmaps(b.w1, b.w2, C, &b.tm_w1, &b.tm_w2));
        b.ffn_fused = true;
      }
    }
  }
  const int Cl = ds.channels[d->nst - 1], V = ds.vae_dim;
  d->w_head_gemm = d->take<bf16>(static_cast<size_t>(V) * Cl * k);        // [dim, Cl, 7] -> [dim, 7*Cl]
  if (!d->w_head_gemm) return VV_ERR_CUDA;
  k_pack_stem_w<T><<<1024, 256, 0, st>>>(static_cast<const T*>(w[wi++]), d->w_head_gemm, V, Cl, k);
  d->b_head_vec = f32vec(w[wi++], V);
  if (!d->b_stem || !d->b_head_vec) return VV_ERR_CUDA;
  d->head_off = off; off += static_cast<size_t>(HIST) * Cl;
  d->slot_elems = (off + 7) / 8 * 8;
  VV_COUNT_SYNC();
  VV_CHECK_CUDA(cudaStreamSynchronize(st));
  return VV_OK;
}

// One pass of the encoder over frames [toff, toff+Tn) (hop samples each) of n utterances.
// TokenizerEncoder.forward (:810-813) / forward_features (:774-808); out [n, Ttot, dim] row-major.
// valid_samples >= 0 (prefill): each utterance holds that many samples (row stride of `audio`), the last frame is
// completed with zeros stage by stage exactly like the non-streaming SConv1d does (:393-401).
int enc_run(VVDec* d, VVCache* c, const long long* slots, const void* audio, int n, int Tn, int Ttot, int toff,
            void* out, int io_dtype, cudaStream_t st, long long valid_samples = -1) {
  CacheView cv;
  cv.base = c->arena.as<bf16>();
  cv.parity = c->parity.as<int>();
  cv.slots = slots;
  cv.copy_stride = static_cast<size_t>(c->phys_slots) * d->slot_elems;
  cv.slot_elems = d->slot_elems;
  bf16* xa = d->xa.as<bf16>();
  bf16* xb = d->xb.as<bf16>();
  bf16* yb = d->ybuf.as<bf16>();
  bf16* hid = d->hid.as<bf16>();
  int Tc = Tn * d->hop;          // samples (rows) per utterance at the current stage
  int rows = n * Tc;
  const bool masked = valid_samples >= 0;
  long long Lg = valid_samples;                              // valid time steps per utterance at the current stage
  long long start = static_cast<long long>(toff) * d->hop;   // first time step of this call at the current stage
  {
    const int C0 = d->st[0].C;
    const size_t total = static_cast<size_t>(rows) * (C0 / 8);
    const size_t Ltot = masked ? static_cast<size_t>(valid_samples) : static_cast<size_t>(Ttot) * d->hop;
    const size_t loff = static_cast<size_t>(toff) * d->hop;
    const int avalid = masked ? static_cast<int>(std::min<long long>(Lg - start, INT_MAX)) : INT_MAX;
    if (io_dtype == VV_DTYPE_BF16)
      VV_CHECK_CUDA(vv_launch(k_enc_stem<bf16>, g1(total, 256), dim3(256), 0, st, 1, true, static_cast<const bf16*>(audio),
                              static_cast<const float*>(d->w_stem_f32), static_cast<const float*>(d->b_stem), xa, cv,
                              d->stem_off, n, Tc, C0, Ltot, loff, avalid));
    else
      VV_CHECK_CUDA(vv_launch(k_enc_stem<float>, g1(total, 256), dim3(256), 0, st, 1, true, static_cast<const float*>(audio),
                              static_cast<const float*>(d->w_stem_f32), static_cast<const float*>(d->b_stem), xa, cv,
                              d->stem_off, n, Tc, C0, Ltot, loff, avalid));
  }
  for (int i = 0; i < d->nst; ++i) {
    const StageW& s = d->st[i];
    if (i > 0) {
      // ---- strided SConv1d (k = 2*stride) as one GEMM over [block[t] | block[t-1]] ----
      const int Ci = d->st[i - 1].C, r = s.ratio, W = r * Ci;
      rows /= r; Tc /= r;
      const int valid_t = masked ? static_cast<int>(std::min<long long>(Lg - start, INT_MAX)) : INT_MAX;
      Lg = (Lg + r - 1) / r; start /= r;
      VV_CHECK_CUDA(vv_launch(k_convtr_pack, g1(static_cast<size_t>(rows) * W / 8, 256), dim3(256), 0, st, 1, true,
                              static_cast<const bf16*>(xa), hid, cv, s.up_cache_off, n, Tc, W, valid_t, Ci));
      GemmArgs a;
      a.A = hid; a.lda = 2 * W; a.B = s.w_up; a.ldb = 2 * W; a.M = rows; a.N = s.C; a.K = 2 * W;
      a.mode = GM_STORE_BF16; a.out = xb; a.ldo = a.N; a.bias = s.b_up;
      VV_TRY(gemm_bias_splitk(d, a, st));
      bf16* t = xa; xa = xb; xb = t;
    }
    VV_TRY(run_stage_blocks(d, s, cv, n, Tc, rows, xa, xb, yb, hid, st));
  }
  // ---- head conv k=7, C_last -> dim (self.norm is Identity, :767-770): im2col + GEMM ----
  {
    const int Cl = d->st[d->nst - 1].C, V = d->V;
    const int K = (HIST + 1) * Cl;
    const size_t total = static_cast<size_t>(rows) * K + static_cast<size_t>(n) * HIST * Cl;
    VV_CHECK_CUDA(vv_launch(k_stem_pack<bf16>, g1(total, 256), dim3(256), 0, st, 1, true, static_cast<const bf16*>(xa), 1,
                            1.0f, 0.0f, hid, cv, d->head_off, n, Tc, Cl, Tc, 0));
    GemmArgs a;
    a.A = hid; a.lda = K; a.B = d->w_head_gemm; a.ldb = K; a.M = rows; a.N = V; a.K = K;
    const size_t esz = (io_dtype == VV_DTYPE_BF16) ? 2 : 4;
    const bool direct = (Tn == Ttot);            // a time chunk of [n, Ttot, dim] is not contiguous over n
    a.bias = d->b_head_vec; a.ldo = V; a.out = direct ? out : static_cast<void*>(yb);
    a.mode = (io_dtype == VV_DTYPE_BF16) ? GM_STORE_BF16 : GM_STORE_F32;
    VV_TRY(gemm_bias_splitk(d, a, st));
    if (!direct)
      VV_CHECK_CUDA(cudaMemcpy2DAsync(static_cast<char*>(out) + static_cast<size_t>(toff) * V * esz,
                                      static_cast<size_t>(Ttot) * V * esz, yb, static_cast<size_t>(Tn) * V * esz,
                                      static_cast<size_t>(Tn) * V * esz, n, cudaMemcpyDeviceToDevice, st));
  }
  VV_CHECK_CUDA(vv_launch(k_cache_flip, g1(n, 128), dim3(128), 0, st, 1, true, c->parity.as<int>(), slots, n));
  VV_CHECK_CUDA(cudaGetLastError());
  return VV_OK;
}

// Non-streaming encode (:774-813 with cache=None): the streaming path on a private zeroed cache, chunked in time
// so that scratch stays bounded by max_rows frames; zero history == the causal left zero padding (:398-401).
int enc_offline(VVDec* e, const void* audio, int n, int T, long long valid_samples, void* out, int io_dtype,
                cudaStream_t st, const char* who) {
  VV_REQUIRE(n <= e->desc.max_rows, "%s: n=%d exceeds max_rows=%d", who, n, e->desc.max_rows);
  if (e->priv == nullptr || e->priv->max_slots < n) {
    VV_COUNT_SYNC();
    VV_CHECK_CUDA(cudaStreamSynchronize(st));
    delete e->priv;
    e->priv = nullptr;
    VV_TRY(cache_alloc(e, n, &e->priv));
  }
  k_cache_zero<<<256, 256, 0, st>>>(e->priv->arena.as<bf16>(), nullptr, n,
                                    static_cast<size_t>(e->priv->phys_slots) * e->slot_elems, e->slot_elems);
  VV_COUNT_LAUNCH();
  const int Tc = e->desc.max_rows / n;
  for (int t0 = 0; t0 < T; t0 += Tc) {
    const int Tn = (T - t0 < Tc) ? (T - t0) : Tc;
    VV_TRY(enc_run(e, e->priv, e->priv->iota.as<long long>(), audio, n, Tn, T, t0, out, io_dtype, st, valid_samples));
  }
  return VV_OK;
}

}  // namespace

extern "C" {

int vv_enc_create(const VVDecDesc* desc, const void* const* weights, int n_weights, VVDec** out) {
  VV_REQUIRE(desc && weights && out, "vv_enc_create: null argument");
  const int nst = desc->n_stages;
  VV_REQUIRE(nst >= 1 && nst <= VV_DEC_MAX_STAGES, "vv_enc_create: n_stages=%d", nst);
  VV_REQUIRE(desc->kernel == HIST + 1, "vv_enc_create: kernel=%d (only 7 is supported)", desc->kernel);
  VV_REQUIRE(desc->vae_dim > 0 && desc->vae_dim % 8 == 0, "vv_enc_create: dim=%d must be a multiple of 8", desc->vae_dim);
  VV_REQUIRE(desc->max_rows >= 1, "vv_enc_create: max_rows=%d", desc->max_rows);
  int expect = 2 + 2;
  int hop = 1;
  for (int i = 1; i < nst; ++i) {
    VV_REQUIRE(desc->ratios[i] >= 1 && desc->ratios[i] <= 64, "vv_enc_create: ratios[%d]=%d", i, desc->ratios[i]);
    hop *= desc->ratios[i];
  }
  size_t row_elems = 0;
  int down = 1;
  for (int i = 0; i < nst; ++i) {
    const int C = desc->channels[i];
    VV_REQUIRE(C >= 8 && C <= 2048 && (C & (C - 1)) == 0, "vv_enc_create: channels[%d]=%d must be a power of two in [8,2048]", i, C);
    VV_REQUIRE(desc->depths[i] >= 0 && desc->depths[i] <= 64, "vv_enc_create: depths[%d]=%d", i, desc->depths[i]);
    if (i > 0) { down *= desc->ratios[i]; expect += 2; }
    expect += 10 * desc->depths[i];
    const size_t e = static_cast<size_t>(hop / down) * C;     // elements per frame at this stage
    if (e > row_elems) row_elems = e;
  }
  VV_REQUIRE(n_weights == expect, "vv_enc_create: expected %d weights, got %d", expect, n_weights);
  for (int i = 0; i < n_weights; ++i) VV_REQUIRE(weights[i] != nullptr, "vv_enc_create: weight %d is null", i);
  VV_REQUIRE(desc->weight_dtype == VV_DTYPE_BF16 || desc->weight_dtype == VV_DTYPE_F32, "vv_enc_create: weight_dtype");
  VVDec* d = new VVDec();
  d->encoder = true;
  d->desc = *desc; d->V = desc->vae_dim; d->nst = nst; d->hop = hop; d->row_elems = row_elems;
  int rc = (desc->weight_dtype == VV_DTYPE_F32) ? pack_all_enc<float>(d, weights, 0) : pack_all_enc<bf16>(d, weights, 0);
  const size_t R = desc->max_rows;
  // hid holds the 4C-wide FFN hidden, the strided-conv operand (2x its input) and the head im2col (7 x C_last)
  size_t hid_elems = 4 * row_elems;
  const size_t head_elems = static_cast<size_t>(HIST + 1) * desc->channels[nst - 1];
  if (head_elems > hid_elems) hid_elems = head_elems;
  if (rc == VV_OK) rc = d->xa.alloc(R * row_elems * 2);
  if (rc == VV_OK) rc = d->xb.alloc(R * row_elems * 2);
  if (rc == VV_OK) rc = d->ybuf.alloc(R * row_elems * 2);
  if (rc == VV_OK) rc = d->hid.alloc(R * hid_elems * 2);
  d->partial_elems = dec_partial_elems(R);
  if (rc == VV_OK) rc = d->partial.alloc(d->partial_elems * sizeof(float));
  if (rc == VV_OK) {
    cudaError_t e = cudaFuncSetAttribute(k_block_dw, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_block_dw_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) {
      vv_set_error("vv_enc_create: %s", cudaGetErrorString(e));
      rc = VV_ERR_CUDA;
    }
  }
  if (rc != VV_OK) { delete d; return rc; }
  *out = d;
  return VV_OK;
}

int vv_enc_step(VVDec* e, VVCache* c, const void* audio, const int64_t* slot_idx, int n, int T, void* out, int io_dtype,
                void* stream) {
  VV_REQUIRE(e && c && audio && slot_idx && out, "vv_enc_step: null argument");
  VV_REQUIRE(e->encoder, "vv_enc_step: handle is a decoder");
  VV_REQUIRE(c->d == e, "vv_enc_step: cache belongs to another handle");
  VV_REQUIRE(n >= 1 && T >= 1, "vv_enc_step: n=%d T=%d", n, T);
  VV_REQUIRE(n <= c->max_slots, "vv_enc_step: n=%d exceeds cache slots %d", n, c->max_slots);
  VV_REQUIRE(static_cast<long long>(n) * T <= e->desc.max_rows, "vv_enc_step: n*T=%lld exceeds max_rows=%d",
             static_cast<long long>(n) * T, e->desc.max_rows);
  VV_REQUIRE(io_dtype == VV_DTYPE_BF16 || io_dtype == VV_DTYPE_F32, "vv_enc_step: io_dtype");
  const long long* slots = nullptr;
  VV_TRY(cache_check_slots(c, reinterpret_cast<const long long*>(slot_idx), n, static_cast<cudaStream_t>(stream), &slots,
                           "vv_enc_step"));
  VV_TRY(enc_run(e, c, slots, audio, n, T, T, 0, out, io_dtype, static_cast<cudaStream_t>(stream)));
  return cache_publish_errors(c, static_cast<cudaStream_t>(stream));
}

int vv_enc_encode(VVDec* e, const void* audio, int n, int T, void* out, int io_dtype, void* stream) {
  VV_REQUIRE(e && audio && out, "vv_enc_encode: null argument");
  VV_REQUIRE(e->encoder, "vv_enc_encode: handle is a decoder");
  VV_REQUIRE(n >= 1 && T >= 1, "vv_enc_encode: n=%d T=%d", n, T);
  VV_REQUIRE(io_dtype == VV_DTYPE_BF16 || io_dtype == VV_DTYPE_F32, "vv_enc_encode: io_dtype");
  return enc_offline(e, audio, n, T, -1, out, io_dtype, static_cast<cudaStream_t>(stream), "vv_enc_encode");
}

int vv_enc_prefill(VVDec* e, const void* audio, int n, long long n_samples, void* out, int io_dtype, void* stream) {
  VV_REQUIRE(e && audio && out, "vv_enc_prefill: null argument");
  VV_REQUIRE(e->encoder, "vv_enc_prefill: handle is a decoder");
  VV_REQUIRE(n >= 1 && n_samples >= 1, "vv_enc_prefill: n=%d n_samples=%lld", n, n_samples);
  VV_REQUIRE(n_samples <= static_cast<long long>(INT_MAX) / 2, "vv_enc_prefill: n_samples=%lld too long", n_samples);
  VV_REQUIRE(io_dtype == VV_DTYPE_BF16 || io_dtype == VV_DTYPE_F32, "vv_enc_prefill: io_dtype");
  const int T = static_cast<int>((n_samples + e->hop - 1) / e->hop);
  return enc_offline(e, audio, n, T, n_samples, out, io_dtype, static_cast<cudaStream_t>(stream), "vv_enc_prefill");
}

void vv_enc_destroy(VVDec* e) { delete e; }

}  // extern "C"
