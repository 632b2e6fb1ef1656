// vv_gemm.cuh — the one tensor-core contraction every dense op of the path uses.
//
//   D[M,N] = epilogue( A[M,K] · B[N,K]^T )     A, B bf16 row-major (K contiguous)
//
// B is exactly the nn.Linear weight layout, so the head's Linear layers
// (modular_vibevoice_diffusion_head.py:59-62,111-113,155,177,181,213-214), the decoder's
// pointwise FFN (modular_vibevoice_tokenizer.py:588-596), the dense stem conv (im2col,
// K = 7*64) and the transposed convs (two taps => K = 2*Cin, N = stride*Cout) all map here.
//
// B200 mapping: TMA (SWIZZLE_128B boxes of 64 bf16 along K) -> shared-memory ring ->
// tcgen05.mma.cta_group::1.kind::f16 (UMMA 128 x BN x 16) issued by one thread ->
// fp32 accumulator in TMEM -> 8 epilogue warps (tcgen05.ld 32x32b.x32) apply the fused
// epilogue and store with 16-byte vectors.  Rows/cols/K beyond the problem are zero-filled
// by TMA on load and masked on store, so M < 128 (streaming batches) needs no padding.
#pragma once
#include "vv_common.cuh"

enum GemmMode : int {
  GM_STORE_BF16 = 0,        // out_bf16 = acc + bias
  GM_GELU_BF16 = 1,         // out_bf16 = gelu_erf(acc + bias)
  GM_SWIGLU_BF16 = 2,       // B rows interleaved (gate_j, up_j); out_bf16[m, n/2] = silu(g)*u
  GM_RESID_GAMMA_BF16 = 3,  // out_bf16 = resid_bf16 + gamma[n] * (acc + bias[n])
  GM_PARTIAL_F32 = 4,       // split-K partial: out_f32[z][m][n] = acc
  GM_RESID_GATE_F32 = 5,    // out_f32 = resid_f32 + gate_bf16[m,n] * acc
  GM_STORE_F32 = 6,         // out_f32 = acc + bias
  GM_SILU_BF16 = 7,         // out_bf16 = silu(acc + bias)
};

struct GemmArgs {
  const bf16* A = nullptr; int lda = 0;
  const bf16* B = nullptr; int ldb = 0;
  int M = 0, N = 0, K = 0;
  int mode = GM_STORE_BF16;
  void* out = nullptr; int ldo = 0;
  const float* bias = nullptr;
  const float* gamma = nullptr;
  const void* resid = nullptr; int ldr = 0;
  const bf16* gate = nullptr; int ldg = 0;
  int bn = 0;       // 0 = choose
  int splits = 0;   // 0 = choose (only GM_PARTIAL_F32 may split K)
  int b_streaming = 0;  // 1: B is read once (weight streaming) -> L2 evict-first
  int cluster = 0;  // 0 = choose; CTAs per cluster along N sharing (multicasting) the A tile: 1, 2, 4
  int stages = 0;   // 0 = choose; depth of the TMA -> MMA shared-memory ring
  int pdl = 1;      // launch with programmatic stream serialization (see vv_launch)
  int swap = 0;     // 0 / -1: never; 2: let the planner choose; 1: weights on the MMA's M side, activations (M <= bn rows) on its N side; -1: never
  int persistent = 0;  // 0 = choose; 1: one CTA per SM walks the tiles (double-buffered TMEM accumulator); -1: never
};

// Plans tile width / split count for a problem; fills args.bn / args.splits if 0.
void gemm_plan(GemmArgs& a, int num_sms);
// Launches on `stream`. Returns VV_OK or a negative VVStatus (vv_last_error() set).
int gemm_launch(const GemmArgs& a, cudaStream_t stream);
// Number of fp32 elements of split-K scratch a planned GEMM needs (splits*M*ldo).
size_t gemm_partial_elems(const GemmArgs& a);

// Optional per-launch timing (bench.py's live roofline): when enabled every gemm_launch is
// bracketed by CUDA events on its own stream; gemm_timing_fetch synchronises and returns them.
struct GemmLaunchRec { int M, N, K, bn, splits, mode; float ms; int kind; long long bytes; };
bool gemm_timing_on();
// brackets for other kernels (the persistent head launch): events are created/recorded by the caller
void gemm_timing_push(const GemmLaunchRec& rec, cudaEvent_t e0, cudaEvent_t e1);
void gemm_timing_enable(bool on);
int gemm_timing_fetch(GemmLaunchRec* out, int cap);   // returns number of records written; clears

// Debug builds (make TRACE=1) only: copies the per-CTA clock stamps [160][48]; returns elements written (0 if not a trace build).
int gemm_debug_trace(long long* out, int cap);

// K-major SWIZZLE_128B tensor map of a bf16 matrix [rows, cols] (row stride ld elements) with boxes of
// [box_rows, 64 cols] — the operand layout tcgen05.mma reads.
int gemm_make_tmap(CUtensorMap* m, const bf16* ptr, int rows, int cols, int ld, int box_rows);
