// vv_block.cuh — shared decoder types + the fused Block1D kernel's interface.
#pragma once
#include "vv_common.cuh"

constexpr int HIST = 6;   // kernel_size - 1 rows of history for the k=7 causal convs (tokenizer.py:282)

// Streaming conv cache arena: [2 copies][max_slots][slot_elems] bf16 + a parity bit per slot
// (see vv_dec.cu).  A step reads copy parity[s] and writes copy parity[s]^1.
struct CacheView {
  bf16* base;              // arena
  const int* parity;       // [max_slots]
  const long long* slots;  // [n] cache row of each batch row (device int64)
  size_t copy_stride;      // max_slots * slot_elems
  size_t slot_elems;
};
#ifdef __CUDACC__
__device__ __forceinline__ const bf16* cache_rd(const CacheView& c, int b, size_t off) {
  const long long s = c.slots[b];
  return c.base + static_cast<size_t>(c.parity[s]) * c.copy_stride + static_cast<size_t>(s) * c.slot_elems + off;
}
__device__ __forceinline__ bf16* cache_wr(const CacheView& c, int b, size_t off) {
  const long long s = c.slots[b];
  return c.base + static_cast<size_t>(c.parity[s] ^ 1) * c.copy_stride + static_cast<size_t>(s) * c.slot_elems + off;
}
#endif

// One whole Block1D (modular_vibevoice_tokenizer.py:620-684, inlined at :924-944) in one launch:
//   xn = ConvRMSNorm(x); x' = x + gamma*(dwconv7_causal(xn)+b); y = ConvRMSNorm_ffn(x')
//   x'' = x' + ffn_gamma * (W2 . gelu(W1 . y + b1) + b2)
// for channel-last activations [n, Tn, C] bf16, C in {32, 64, 128}.
struct BlockFusedArgs {
  const bf16* x_in; bf16* x_out;
  const float* norm_w; const float* dw_w;   // dw_w: [7][C] tap-major
  const float* dw_b; const float* gamma; const float* ffn_norm_w;
  const float* b1; const float* b2; const float* ffn_gamma;
  const CUtensorMap* tm_w1;   // W1 [4C, C] bf16, box [128 rows, 64 cols], SWIZZLE_128B
  const CUtensorMap* tm_w2;   // W2 [C, 4C] bf16, box [C rows, 64 cols],  SWIZZLE_128B
  CacheView cv; size_t cache_off;
  int n, Tn, C;
  float eps;
};

bool block_fused_supported(int C, int Tn);
// tensor maps for a block's FFN weights (encoded once at vv_dec_create)
int block_fused_make_tmaps(const bf16* w1, const bf16* w2, int C, CUtensorMap* tm_w1, CUtensorMap* tm_w2);
int block_fused_launch(const BlockFusedArgs& a, cudaStream_t stream);
// Pipelined version of the same block (vv_blockp.cu: mixer of tile i+1 overlaps the GELU / epilogue of tile i); same
// arguments and tensor maps.  block_fused_launch dispatches to it unless vv_set_option("block_pipe", 0).
int block_pipe_launch(const BlockFusedArgs& a, cudaStream_t stream);
int pipe_debug_trace(long long* out, int cap);    // trace builds only
int block_debug_trace(long long* out, int cap);   // trace builds only (make TRACE=1)
int ffn_debug_trace(long long* out, int cap);     // trace builds only

// FFN half of a Block1D for C = 256 in one launch (hidden activations stay on the SM):
//   x_out = xres + ffn_gamma * (W2 . gelu(W1 . y + b1) + b2),   y / xres / x_out: [rows, C] bf16
struct FfnFusedArgs {
  const bf16* y; const bf16* xres; bf16* x_out;
  const float* b1; const float* b2; const float* ffn_gamma;
  const CUtensorMap* tm_w1; const CUtensorMap* tm_w2;
  int rows, C;
};
bool ffn_fused_supported(int C);
int ffn_fused_make_tmaps(const bf16* w1, const bf16* w2, int C, CUtensorMap* tm_w1, CUtensorMap* tm_w2);
int ffn_fused_launch(const FfnFusedArgs& a, cudaStream_t stream);
