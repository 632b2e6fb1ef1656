// vv_chain.cuh — persistent phase engine: a whole sequence of dependent small-M GEMMs and row-wise
// ops (one solver loop of the diffusion head) in ONE cooperative launch.
//
// Why (profiles/r01_gemm_trace.md): at M <= 128 every GEMM launch pays ~3 us of prologue, A-tile
// latency and a cold-instruction-cache epilogue around a 4-7 us weight stream, and a frame of the
// head is ~150 such launches.  Here one CTA per SM stays resident for all S solver steps:
//   * the op list ("program") lives in device memory; every CTA walks it phase by phase;
//   * a GEMM phase hands (n-tile, k-split) work items to CTAs round-robin; the TMA -> tcgen05
//     ring and its mbarrier phases simply continue across phases;
//   * weights never depend on activations, so the producer warp fills the ring with the NEXT
//     GEMM phase's weight tiles while the grid barrier of the current phase is still pending;
//     only the activation (A) tiles wait for the barrier;
//   * the accumulator is double-buffered in TMEM: the MMA of work item i+1 overlaps the epilogue of i;
//   * row-wise phases (split-K reduce + gated residual + RMSNorm + adaLN modulate; CFG combine +
//     DPM-Solver++ update) run on the eight epilogue warps, rows spread over the CTAs;
//   * phases are separated by a grid barrier (arrival counter + a private flag line per CTA).
#pragma once
#include "vv_block.cuh"
#include "vv_gemm.cuh"

struct ChainProgram;   // opaque (vv_chain.cu)

struct ChainNormMod {     // k_norm_mod's arguments (vv_head.cu)
  // rowmajor != 0: the partials lie [M][nsplit][H] (chain_add_gemm's part_rowmajor) - a row's partials are one
  // contiguous block; else [nsplit][M][H]
  float* x; const float* partial; int nsplit; int M; int rowmajor;
  const bf16* prev_gate; const float* w; const bf16* shift; const bf16* scale; int ld_mod;
  bf16* y; int H; float eps;
};
struct ChainCfg {         // k_cfg_update's arguments that are fixed per program
  const float* v; int nsplit; float* zstate; float* prev_x0; bf16* zb; int step;
  // this step's DPM-Solver++ scalars (StepCoef, vv_head.cu): x0 = alpha_s*x - sigma_s*v;
  // x' = cx*x + c0*x0 + c1*(x0 - x0_prev)*inv_r0 (+ cn*noise)
  float alpha_s, sigma_s, cx, c0, c1, inv_r0, cn; int order;
  int last;               // 1: also writes the caller's output latent
  int n, L;
};
// One row-pair phase for the whole step boundary (final layer -> CFG -> solver update -> input
// projection -> first layer's norm), so a solver step costs one grid barrier there instead of five:
//   x += gate * sum_z partial_z ; yF = bf16(rmsnorm(x) * (1 + fscale) + fshift)       (FinalLayer, :184-188)
//   v = yF . w_out^T            (both rows of the pair: cond row p, uncond row p + n)
//   z' = solver(z, uncond + cfg * (cond - uncond))                                     (k_cfg_update)
//   x = bf16(z') . w_in^T  (same for both rows) ; y = bf16(rmsnorm(x) * w0 * (1 + scale0) + shift0)
// `last` stops after the solver update and writes the caller's latent.  L must be 64.
struct ChainTail {
  float* x; const float* partial; const bf16* prev_gate; const bf16* fshift;   // fscale = fshift + H
  const bf16* w_out; float* zstate; float* prev_x0; bf16* zb;
  const bf16* w_in; const float* norm_w0; const bf16* shift0;                  // scale0 = shift0 + H (next step's table)
  bf16* y;
  int nsplit, n, ld_mod, step, last, L, H, order;
  float alpha_s, sigma_s, cx, c0, c1, inv_r0, cn, eps;
  // split != 0: a row pair is shared by TWO CTAs, each owning half of the H columns (its half of the folded rows, of
  // the final Linear's K range and of the input projection's outputs); the two meet three times through `mail`
  // (sums of squares, the final Linear's partial sums, the next norm's sum of squares).  Set by chain_finalize.
  int split;
  float* mail;
  int rowmajor;            // layout of `partial`, see ChainNormMod
};
// Mixer half of a Block1D for ONE new time step per utterance (streaming frame-step, wide stages: rows = utterances
// <= 128, C <= 2048), k_block_dw's arithmetic in the same order (vv_dec.cu):
//   x  = x_in[row]                                   or, when part != null (the previous block's FFN ended in split-K
//        xres[row] + gamma2 * (sum_z part[z][row] + bias2)    partials that nobody reduced), the reduced value, bf16
//   xn = bf16(rmsnorm(x) * norm_w) ; history h[0..5] = the cache's six normalised rows (ChainCall::cv, offset `off`)
//   x' = bf16(x + gamma * (dw_b + sum_j dw_w[j] * [h ; xn][j]))  -> x_out ;  y = bf16(rmsnorm(x') * ffn_norm_w) -> y
//   new history = [h[1..5] ; xn]
struct ChainDw {
  const bf16* x_in; bf16* x_out; bf16* y;
  const float* norm_w; const float* dw_w; const float* dw_b; const float* gamma; const float* ffn_norm_w;
  const float* part; const bf16* xres; const float* gamma2; const float* bias2;
  size_t off;
  int nsplit, rows, C;
  float eps;
};
// Mixer half of a Block1D over Tn new time steps per utterance (wide stages whose rows no longer fit one 128-row tile:
// C = 1024 / 512 at 8 / 40 steps per utterance), k_block_dw's arithmetic in the same order (vv_dec.cu).  A work item
// is one utterance x one tile of TT time steps (+ the 6-row causal halo, re-normalised from x_in or read from the
// cache when it precedes this call); the normalised rows live in the activation ring's staging area.
//   xn = bf16(rmsnorm(x) * norm_w) ; x' = bf16(x + gamma * (dw_b + sum_j dw_w[j] * xn[t-6+j])) -> x_out
//   y = bf16(rmsnorm(x') * ffn_norm_w) -> y ; new history = xn[Tn-6 .. Tn-1]
struct ChainDwTile {
  const bf16* x_in; bf16* x_out; bf16* y;
  const float* norm_w; const float* dw_w; const float* dw_b; const float* gamma; const float* ffn_norm_w;
  size_t off;
  int n, Tn, C, TT;
  float eps;
};
struct ChainCall {        // per-launch values
  float cfg;
  const float* sde_noise; // [S, n, L] or null
  void* out; int out_dtype;
  CacheView cv;           // streaming conv cache of this call (ChainDw phases)
};

ChainProgram* chain_create();
void chain_destroy(ChainProgram* p);
void chain_clear(ChainProgram* p);
// GEMM phase: any M (ceil(M / 128) row tiles; work items = N tiles x row tiles x splits), BN = 64; a.mode in
// {GM_STORE_F32, GM_SWIGLU_BF16, GM_PARTIAL_F32, GM_STORE_BF16, GM_SILU_BF16, GM_GELU_BF16, GM_RESID_GAMMA_BF16};
// a.splits > 1 only with GM_PARTIAL_F32.  Returns VV_OK or an error.
// part_rowmajor (GM_PARTIAL_F32 only): the splits' outputs are stored [M][splits][N] (a.ldo = N) instead of
// [splits][M][ldo], so that the row phase which sums them fetches a row's partials with ONE bulk copy.
int chain_add_gemm(ChainProgram* p, const GemmArgs& a, int part_rowmajor = 0);
int chain_add_norm_mod(ChainProgram* p, const ChainNormMod& a);
int chain_add_cfg(ChainProgram* p, const ChainCfg& a);
int chain_add_tail(ChainProgram* p, const ChainTail& a);
int chain_add_dwmix(ChainProgram* p, const ChainDw& a);
int chain_add_dwtile(ChainProgram* p, const ChainDwTile& a);
void chain_set_kind(ChainProgram* p, int kind);            // kind of the timing record a launch pushes (1: head, 2: decoder/encoder stage)
int chain_finalize(ChainProgram* p);                       // uploads the program (synchronises)
int chain_launch(ChainProgram* p, const ChainCall& call, cudaStream_t stream);
int chain_num_phases(const ChainProgram* p);
int chain_debug_trace(long long* out, int cap);   // trace builds only
