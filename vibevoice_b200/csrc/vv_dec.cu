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
// Streaming cache: ONE arena [2][max_slots][slot_elems] bf16, updated in place by the layer
// kernels.  Each layer keeps exactly the history it needs (6 rows for k=7 convs, 1 row for a
// transposed conv — the reference keeps k-1 = 15/9/7/3 input frames and recomputes them,
// :447,:539-547).  Two copies + a per-slot parity bit make a step race-free: a layer's first
// tile reads history while its last tile writes the new one.  All-zero history == "no
// history" == set_to_zero (:234-241).  Offline decode is the same code path run on a private
// zeroed cache, chunked in time, so there is a single implementation to validate.
#include <vector>

#include "../../include/vvb200.h"
#include "vv_common.cuh"
#include "vv_gemm.cuh"

namespace {

constexpr int HIST = 6;   // kernel_size - 1 rows of history for the k=7 causal convs (:282)

struct CacheView {
  bf16* base;              // arena
  const int* parity;       // [max_slots]
  const long long* slots;  // [n] cache row of each batch row (device int64)
  size_t copy_stride;      // max_slots * slot_elems
  size_t slot_elems;
};
__device__ __forceinline__ const bf16* cache_rd(const CacheView& c, int b, size_t off) {
  const long long s = c.slots[b];
  return c.base + static_cast<size_t>(c.parity[s]) * c.copy_stride + static_cast<size_t>(s) * c.slot_elems + off;
}
__device__ __forceinline__ bf16* cache_wr(const CacheView& c, int b, size_t off) {
  const long long s = c.slots[b];
  return c.base + static_cast<size_t>(c.parity[s] ^ 1) * c.copy_stride + static_cast<size_t>(s) * c.slot_elems + off;
}

// ---------------------------------------------------------------------------------------
// Stem: latent de-normalisation (modeling_vibevoice_inference.py:651) + im2col for the causal
// k=7 dense conv (SConv1d, :327-382 streaming / :384-418 non-streaming).
//   A[(b,t), j*V + ci] = z[b, t-6+j, ci]   (history rows from the cache)
// ---------------------------------------------------------------------------------------
template <typename T>
__global__ void k_stem_pack(const T* __restrict__ lat, int time_major, float inv_scale, float bias,
                            bf16* __restrict__ A, CacheView cv, size_t off, int n, int Tn, int V) {
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
    const size_t src = time_major ? (static_cast<size_t>(b) * Tn + tsrc) * V + ci
                                  : (static_cast<size_t>(b) * V + ci) * Tn + tsrc;
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
                              size_t off, int n, int Tn, int C) {
  const int vpr = C / 8;                               // uint4 vectors per row
  const size_t total = static_cast<size_t>(n) * Tn * vpr;
  const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  if (i >= total) return;
  const int c8 = static_cast<int>(i % vpr);
  const size_t row = i / vpr;
  const int b = static_cast<int>(row / Tn), t = static_cast<int>(row % Tn);
  const uint4 cur = *reinterpret_cast<const uint4*>(x + row * C + c8 * 8);
  uint4 prev;
  if (t > 0) prev = *reinterpret_cast<const uint4*>(x + (row - 1) * C + c8 * 8);
  else prev = *reinterpret_cast<const uint4*>(cache_rd(cv, b, off) + c8 * 8);
  bf16* dst = A2 + row * (2 * static_cast<size_t>(C));
  *reinterpret_cast<uint4*>(dst + c8 * 8) = cur;
  *reinterpret_cast<uint4*>(dst + C + c8 * 8) = prev;
  if (t == Tn - 1) *reinterpret_cast<uint4*>(cache_wr(cv, b, off) + c8 * 8) = cur;
}

// par[slot] ^= 1 for the rows stepped this call (after every layer wrote its new history).
__global__ void k_cache_flip(int* __restrict__ parity, const long long* __restrict__ slots, int n) {
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
