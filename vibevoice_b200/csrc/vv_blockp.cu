// vv_blockp.cu — pipelined fused Block1D for the narrow decoder / encoder stages (C = 32 / 64 / 128).
//
// Reference: Block1D (modular_vibevoice_tokenizer.py:620-684, inlined in TokenizerDecoder
// .forward_features :924-944), ConvRMSNorm (:77-91), FFN (:579-596), SConv1d streaming (:327-382).
//
// Same arithmetic, in the same order, as k_block_fused (vv_block.cu), but the roles no longer take turns.
// k_block_fused ran mixer -> MMA1 -> GELU -> MMA2 -> epilogue with all 8 compute warps in lock-step, one
// tile in flight per SM (ncu r01: 15.7 % warps active, 39 % long scoreboard, ~9 % of its roofline).  Here
// one persistent CTA per SM keeps TWO tiles in flight:
//   warp 0       weight producer (TMA): C <= 64 the block's W1/W2 are loaded once and stay resident;
//                C = 128 they stream through a ring of 16 KB k-blocks
//   warp 1       MMA issuer: D1[g&1] = Y . W1c^T ; D2[tile&1] += H . W2c^T  (TMEM: 2 x 128 + 2 x C columns)
//   warps 2-17   16 GELU / epilogue warps (4 per scheduler):
//                tcgen05.ld D1 -> +b1 -> exact-erf GELU -> H (UMMA operand layout); then
//                tcgen05.ld D2 -> x'' = x' + ffn_gamma * (D2 + b2) -> global, one tile LATER (after the next
//                tile's GELU chunks), so nobody waits for the last MMA2 of a tile
//   warps 18-25  8 mixer warps working one tile AHEAD of the epilogue, 16 rows each and no barrier between
//                them: RMSNorm in place in the row buffer -> 7-tap depthwise conv, two rows per thread
//                so a loaded row feeds two outputs -> layer-scale residual x' (to global: x_out doubles
//                as the x' buffer, the epilogue reads it back through L1) -> FFN RMSNorm -> Y[tile&1]
//   warp 26      row producer: one bulk copy (cp.async.bulk) per tile into the double-buffered row buffer
// Y is double buffered, so the mixer of tile i+1 overlaps the GELU phases of tile i; H is double buffered
// where shared memory allows (C <= 64).  All per-channel vectors (taps, norms, biases) live in shared
// memory, so no role needs more than the 72 registers that 27 resident warps leave per thread.
#include "vv_block.cuh"

#include "../../include/vvb200.h"
#include "vv_gemm.cuh"

namespace {

constexpr int BLK_M = 128;              // rows (time steps of one utterance) per tile
constexpr int HCH = 128;                // hidden-dim chunk
constexpr int KB_BYTES = BLK_M * 128;   // one [128 x 64] bf16 K-major SWIZZLE_128B tile
constexpr int P_EPI_WARPS = 16;
constexpr int P_MIX_WARPS = 8;              // 2 per scheduler: with 1 the mixer is latency-bound (7.1 k clk per C = 32 tile, r02 trace)
constexpr int P_WARP_EPI0 = 2;
constexpr int P_WARP_MIX0 = P_WARP_EPI0 + P_EPI_WARPS;
constexpr int P_WARP_ROWS = P_WARP_MIX0 + P_MIX_WARPS;         // row producer (bulk copies of the tiles' raw rows)
constexpr int P_THREADS = 32 * (P_WARP_ROWS + 1);             // 864
constexpr int MIX_THREADS = 32 * P_MIX_WARPS;                 // 256
constexpr int TMEM_D2 = 2 * HCH;

constexpr int r16(int v) { return (v + 15) / 16 * 16; }

template <int C>
struct PCfg {
  static constexpr int NCH = 4 * C / HCH;                     // hidden chunks per tile: 1, 2, 4
  static constexpr int NKB1 = (C + 63) / 64;                  // 64-col k-blocks of Y / W1: 1, 1, 2
  static constexpr int NK1 = C / 16;                          // UMMA_K slices of MMA1
  static constexpr bool RESIDENT = (C <= 64);                 // whole W1 / W2 stay in shared memory
  static constexpr int W1_CHUNK = NKB1 * KB_BYTES;            // [128 hidden x C(->64)]
  static constexpr int W2_KB = C * 128;                       // one [C x 64] box
  static constexpr int W2_CHUNK = 2 * W2_KB;
  static constexpr int NH = (C == 128) ? 1 : 2;               // H buffers
  static constexpr int NSTAGE = 3;                            // ring mode (C = 128): 16 KB units
  static constexpr int W_BYTES = RESIDENT ? NCH * (W1_CHUNK + W2_CHUNK) : NSTAGE * KB_BYTES;
  static constexpr int XB_BYTES = r16((BLK_M + HIST) * C * 2);   // one row buffer: the tile's rows incl. halo, as they lie in global memory
  static constexpr int Y_BYTES = NKB1 * KB_BYTES;
  static constexpr int H_BYTES = 2 * KB_BYTES;                // [128 x 128] bf16
  // per-channel vectors (floats): wn[C] taps[7][C] bb[C] gm[C] wf[C] b2[C] fg[C] b1[4C]
  static constexpr int PAR_WN = 0, PAR_TAPS = C, PAR_BB = 8 * C, PAR_GM = 9 * C, PAR_WF = 10 * C, PAR_B2 = 11 * C,
                       PAR_FG = 12 * C, PAR_B1 = 13 * C, PAR_FLOATS = 17 * C;
  static constexpr int OFF_Y = 0;                             // [2]
  static constexpr int OFF_H = OFF_Y + 2 * Y_BYTES;           // [NH]
  static constexpr int OFF_W = OFF_H + NH * H_BYTES;
  static constexpr int OFF_XB = OFF_W + W_BYTES;               // [2]
  static constexpr int OFF_PAR = OFF_XB + 2 * XB_BYTES;
  static constexpr int OFF_BAR = OFF_PAR + PAR_FLOATS * 4;
  static constexpr int SMEM_BYTES = 1024 + OFF_BAR + 256;
  static_assert(SMEM_BYTES <= 227 * 1024, "k_block_pipe shared memory");
  static_assert(OFF_W % 1024 == 0 && OFF_H % 1024 == 0 && OFF_XB % 16 == 0 && OFF_BAR % 8 == 0, "alignment");
};

struct PipeParams {
  const bf16* x_in; bf16* x_out;
  const float* norm_w; const float* dw_w; const float* dw_b; const float* gamma; const float* ffn_norm_w;
  const float* b1; const float* b2; const float* ffn_gamma;
  CacheView cv; size_t cache_off;
  int n, Tn, full_per_utt, n_full, n_tiles;   // tiles: the n_full whole 128-row tiles first, then one ragged tile per utterance
  float eps;
};

#ifdef VV_GEMM_TRACE
__device__ long long g_pipe_trace[160][48];
#ifndef VV_TRACE_C
#define VV_TRACE_C 32     // the launches of this width leave their stamps (the 9 launches of a step share the buffer)
#endif
#define PTRACE(slot, val) do { if (C == VV_TRACE_C && (slot) < 48) g_pipe_trace[blockIdx.x][slot] = (val); } while (0)
#else
#define PTRACE(slot, val) do { } while (0)
#endif

__device__ __forceinline__ void tmem_ld_x8(uint32_t taddr, uint32_t* r) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void tmem_ld_x16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// Address of the 16-byte chunk (row, chunk) inside a [rows x 64] bf16 K-major SWIZZLE_128B tile.
__device__ __forceinline__ uint8_t* sw128(uint8_t* tile, int row, int chunk) {
  return tile + row * 128 + ((chunk ^ (row & 7)) << 4);
}
template <int LPR>
__device__ __forceinline__ float row_sum_lpr(float v) {
#pragma unroll
  for (int o = LPR >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ void lds8(const float* p, float (&f)[8]) {   // 8 floats from shared memory
  const float4 a = *reinterpret_cast<const float4*>(p), b = *(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// 16-byte global load that the compiler may neither hoist nor turn into LDG.NC (x' was written by another
// warp of this CTA during this launch; same SM, so L1 is coherent for it)
__device__ __forceinline__ uint4 ldg_coherent(const void* p) {
  uint4 v;
  asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void tile_coords(const PipeParams& p, int tile, int& b, int& t0) {
  // whole tiles are numbered first so that the short (ragged) tiles land on the CTAs that got one whole tile fewer
  if (tile < p.n_full) { b = tile / p.full_per_utt; t0 = (tile % p.full_per_utt) * BLK_M; }
  else                 { b = tile - p.n_full;       t0 = p.full_per_utt * BLK_M; }
}

__device__ __forceinline__ float2 bf2_f2(uint32_t v) {        // == __bfloat1622float2
  // PRMT on the alu pipe: ptxas turns `v << 16` into IMAD.U32 on the fma pipe, the one this kernel saturates
  uint32_t lo, hi;
  asm("prmt.b32 %0, %1, 0, 0x1044;" : "=r"(lo) : "r"(v));
  asm("prmt.b32 %0, %1, 0, 0x3244;" : "=r"(hi) : "r"(v));
  return make_float2(__uint_as_float(lo), __uint_as_float(hi));
}
// f[i] * a * w[i] for 8 values, each product rounded as in the scalar code: 8 packed instead of 16 scalar multiplies
__device__ __forceinline__ uint4 scale8_pack(const float (&f)[8], float a, const float (&w)[8]) {
  const float2 aa = make_float2(a, a);
  uint32_t o[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 q = __fmul2_rn(__fmul2_rn(make_float2(f[2 * i], f[2 * i + 1]), aa), make_float2(w[2 * i], w[2 * i + 1]));
    o[i] = pack_bf16x2(q.x, q.y);
  }
  return make_uint4(o[0], o[1], o[2], o[3]);
}
__device__ __forceinline__ void unpack8_f2(const uint4& v, float2 (&f)[4]) {
  f[0] = bf2_f2(v.x); f[1] = bf2_f2(v.y); f[2] = bf2_f2(v.z); f[3] = bf2_f2(v.w);
}
__device__ __forceinline__ void lds8_f2(const float* p, float2 (&f)[4]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *(reinterpret_cast<const float4*>(p) + 1);
  f[0] = make_float2(a.x, a.y); f[1] = make_float2(a.z, a.w); f[2] = make_float2(b.x, b.y); f[3] = make_float2(b.z, b.w);
}

// ---------------------------------------------------------------------------------------
// Mixer.  Each of the 8 mixer warps owns WROWS = 16 rows of the tile.  The tile's raw rows (+ 6-row causal halo) are
// bulk-copied by the row-producer warp into one of two row buffers XB[tile & 1]; a warp normalises ITS rows in place
// (warp 0 also the halo rows), signals norm_done[w], waits for norm_done[w-1] (its conv reads the last 6 rows of the
// warp below) and runs the 7-tap conv over the buffer.  There is no barrier across the mixer warps, so they drift
// apart and fill each other's latency gaps (with a CTA-wide bar.sync between the phases every warp sat in the same
// LDS / shuffle latency at the same time: 12 % of the mixer's samples were issuing, ncu r02).
//
// Phase 1 (one warp): ConvRMSNorm in place.  Halo rows that precede this call come from the streaming cache already
// normalised.  8 channels per lane-item, LPR = C/8 consecutive lanes per row.
// ---------------------------------------------------------------------------------------
constexpr int WROWS = BLK_M / P_MIX_WARPS;   // 16

template <int C>
__device__ __forceinline__ void warp_mixer_norm(const PipeParams& p, const uint8_t* smem, uint8_t* xb, int t0, int w,
                                                int rows_here, int lane) {
  using Cfg = PCfg<C>;
  constexpr int LPR = C / 8;
  const float* par = reinterpret_cast<const float*>(smem + Cfg::OFF_PAR);
  const int c0 = (lane % LPR) * 8;
  const int row0 = (w == 0) ? 0 : w * WROWS + HIST;               // first buffer row of this warp (buffer row = tile row + 6)
  const int items = ((w == 0) ? WROWS + HIST : WROWS) * LPR;      // warp-uniform
  float wn[8];
  lds8(par + Cfg::PAR_WN + c0, wn);
#pragma unroll 4
  for (int i0 = 0; i0 < items; i0 += 32) {
    const int item = i0 + lane;
    const int br = row0 + item / LPR;                             // buffer row
    const int r = br - HIST;                                      // tile row (-6 .. 127)
    const bool from_x = item < items && r < rows_here && (t0 + r) >= 0;
    uint8_t* ptr = xb + br * (C * 2) + c0 * 2;
    uint4 hv = make_uint4(0u, 0u, 0u, 0u);
    float f[8];
    float ss = 0.f;
    if (from_x) {
      hv = *reinterpret_cast<const uint4*>(ptr);
      unpack_bf16x8(hv, f);
#pragma unroll
      for (int i = 0; i < 8; ++i) ss = fmaf(f[i], f[i], ss);
    }
    ss = row_sum_lpr<LPR>(ss);
    if (from_x) {
      const float rms = rsqrtf(ss / static_cast<float>(C) + p.eps);
      *reinterpret_cast<uint4*>(ptr) = scale8_pack(f, rms, wn);
    }
  }
}

// ---------------------------------------------------------------------------------------
// Phase 2 (one warp): 7-tap depthwise conv over the row buffer, layer-scale residual x' (-> global x_out, bf16: the stored
// stream), FFN RMSNorm -> Y in the UMMA K-major SWIZZLE_128B operand layout.  Each lane computes two consecutive rows
// for its 8 channels per pass: the eight buffer rows it loads feed both outputs; taps come from shared memory.
// ---------------------------------------------------------------------------------------
template <int C>
__device__ __forceinline__ void warp_mixer_conv(const PipeParams& p, const uint8_t* smem, const uint8_t* xb,
                                                uint8_t* ytile, int b, int t0, int wrow0, int rows_here, int lane) {
  using Cfg = PCfg<C>;
  constexpr int LPR = C / 8;
  constexpr int RS = 32 / LPR;            // row slots per warp
  constexpr int RPP = 2 * RS;
  constexpr int NP2 = WROWS / RPP;        // 1, 2, 4
  const float* par = reinterpret_cast<const float*>(smem + Cfg::OFF_PAR);
  const int rslot = lane / LPR;
  const int c0 = (lane % LPR) * 8;
  const bf16* xin = p.x_in + (static_cast<size_t>(b) * p.Tn + t0) * C + c0;
  bf16* xout = p.x_out + (static_cast<size_t>(b) * p.Tn + t0) * C + c0;
  uint8_t* ykb = ytile + (c0 >> 6) * KB_BYTES;
  const int ychunk = (c0 & 63) >> 3;
#pragma unroll 1
  for (int ps = 0; ps < NP2; ++ps) {
    const int r0 = wrow0 + ps * RPP + 2 * rslot;   // tile row; out row r reads buffer rows r .. r+6
    const bool live0 = r0 < rows_here, live1 = r0 + 1 < rows_here;
    uint4 xr0 = make_uint4(0u, 0u, 0u, 0u), xr1 = xr0;                  // raw x[t] for the residual
    if (live0) xr0 = *reinterpret_cast<const uint4*>(xin + static_cast<size_t>(r0) * C);
    if (live1) xr1 = *reinterpret_cast<const uint4*>(xin + static_cast<size_t>(r0 + 1) * C);
    float2 a0[4], a1[4];
    lds8_f2(par + Cfg::PAR_BB + c0, a0);
#pragma unroll
    for (int i = 0; i < 4; ++i) a1[i] = a0[i];
    if (live0) {
      float2 wprev[4];
#pragma unroll
      for (int j = 0; j <= HIST + 1; ++j) {
        float2 v[4], w[4];
        unpack8_f2(*reinterpret_cast<const uint4*>(xb + (r0 + j) * (C * 2) + c0 * 2), v);
        if (j <= HIST) {
          lds8_f2(par + Cfg::PAR_TAPS + j * C + c0, w);
#pragma unroll
          for (int i = 0; i < 4; ++i) a0[i] = __ffma2_rn(w[i], v[i], a0[i]);
        }
        if (j >= 1) {
#pragma unroll
          for (int i = 0; i < 4; ++i) a1[i] = __ffma2_rn(wprev[i], v[i], a1[i]);
        }
        if (j <= HIST) {
#pragma unroll
          for (int i = 0; i < 4; ++i) wprev[i] = w[i];
        }
      }
    }
    float2 gm[4];
    lds8_f2(par + Cfg::PAR_GM + c0, gm);
    float o0[8], o1[8];
    float ss0 = 0.f, ss1 = 0.f;
    {
      float2 x0[4], x1[4];
      unpack8_f2(xr0, x0);
      unpack8_f2(xr1, x1);
      float t0v[8], t1v[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 q0 = __ffma2_rn(gm[i], a0[i], x0[i]), q1 = __ffma2_rn(gm[i], a1[i], x1[i]);
        t0v[2 * i] = q0.x; t0v[2 * i + 1] = q0.y; t1v[2 * i] = q1.x; t1v[2 * i + 1] = q1.y;
      }
      const uint4 ov0 = pack_bf16x8(t0v), ov1 = pack_bf16x8(t1v);
      if (live0) *reinterpret_cast<uint4*>(xout + static_cast<size_t>(r0) * C) = ov0;
      if (live1) *reinterpret_cast<uint4*>(xout + static_cast<size_t>(r0 + 1) * C) = ov1;
      unpack_bf16x8(ov0, o0);                 // the FFN norm sees the stored (bf16) residual stream
      unpack_bf16x8(ov1, o1);
      if (live0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) ss0 = fmaf(o0[i], o0[i], ss0);
      }
      if (live1) {
#pragma unroll
        for (int i = 0; i < 8; ++i) ss1 = fmaf(o1[i], o1[i], ss1);
      }
    }
    ss0 = row_sum_lpr<LPR>(ss0);
    ss1 = row_sum_lpr<LPR>(ss1);
    float wf[8];
    lds8(par + Cfg::PAR_WF + c0, wf);
    uint4 y0 = make_uint4(0u, 0u, 0u, 0u), y1 = y0;   // rows past the utterance end: zeros (their outputs are masked)
    if (live0) {
      const float rms = rsqrtf(ss0 / static_cast<float>(C) + p.eps);
      y0 = scale8_pack(o0, rms, wf);
    }
    if (live1) {
      const float rms = rsqrtf(ss1 / static_cast<float>(C) + p.eps);
      y1 = scale8_pack(o1, rms, wf);
    }
    *reinterpret_cast<uint4*>(sw128(ykb, r0, ychunk)) = y0;
    *reinterpret_cast<uint4*>(sw128(ykb, r0 + 1, ychunk)) = y1;
  }
}

template <int C>
__global__ void __launch_bounds__(P_THREADS, 1)
k_block_pipe(const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmW2, const PipeParams p) {
  using Cfg = PCfg<C>;
  constexpr int NCH = Cfg::NCH;
  constexpr int NSTAGE = Cfg::NSTAGE;
  extern __shared__ uint8_t smem_raw[];
  // offset arithmetic on the __shared__ array (not an integer round trip): ptxas keeps LDS / STS instead of generic LD / ST
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* wbuf = smem + Cfg::OFF_W;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::OFF_BAR);
  uint64_t* w_full = bars;                  // TMA -> MMA: resident weights landed (C <= 64)
  uint64_t* full_bar = w_full + 1;          // [NSTAGE] TMA -> MMA (C = 128)
  uint64_t* empty_bar = full_bar + NSTAGE;  // [NSTAGE] MMA -> TMA                                  (commit)
  uint64_t* xb_full = empty_bar + NSTAGE;   // [2] bulk copy -> mixer: raw rows of the tile landed in XB[tile & 1]
  uint64_t* xb_free = xb_full + 2;          // [2] mixer -> row producer: every mixer warp is done with XB[.]  (8 warps)
  uint64_t* norm_done = xb_free + 2;        // [2][8] mixer warp w -> warp w+1: rows of warp w are normalised
  uint64_t* y_full = norm_done + 2 * P_MIX_WARPS;         // [2] mixer -> MMA + epilogue: Y in smem, x' in global  (4 warps)
  uint64_t* y_free = y_full + 2;            // [2] MMA (last MMA1 read Y) + epilogue (saw y_full) -> mixer (1 + 16)
  uint64_t* d1_full = y_free + 2;           // [2] MMA -> epilogue: D1 buffer holds chunk g           (commit)
  uint64_t* h_ready = d1_full + 2;          // [2] epilogue -> MMA: H of chunk g in smem, D1 drained  (16 warps)
  uint64_t* h_free = h_ready + 2;           // [2] MMA -> epilogue: MMA2 finished reading H buffer    (commit)
  uint64_t* d2_full = h_free + 2;           // [2] MMA -> epilogue: D2[tile & 1] complete                 (commit)
  uint64_t* d2_free = d2_full + 2;          // [2] epilogue -> MMA: D2[tile & 1] has been read            (16 warps)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(d2_free + 2);
  float* par = reinterpret_cast<float*>(smem + Cfg::OFF_PAR);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmW1);
    tma_prefetch_desc(&tmW2);
    mbar_init(w_full, 1);
    for (int s = 0; s < NSTAGE; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2 * P_MIX_WARPS; ++s) mbar_init(&norm_done[s], 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&xb_full[s], 1);
      mbar_init(&xb_free[s], P_MIX_WARPS);
      mbar_init(&y_full[s], P_MIX_WARPS);
      mbar_init(&y_free[s], 1 + P_EPI_WARPS);
      mbar_init(&d1_full[s], 1);
      mbar_init(&h_ready[s], P_EPI_WARPS);
      mbar_init(&h_free[s], 1);
      mbar_init(&d2_full[s], 1);
      mbar_init(&d2_free[s], P_EPI_WARPS);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  // per-channel vectors -> shared memory (weights: nothing here depends on the predecessor kernel)
  for (int i = threadIdx.x; i < Cfg::PAR_FLOATS; i += P_THREADS) {
    float v;
    if (i < Cfg::PAR_TAPS) v = __ldg(p.norm_w + i);
    else if (i < Cfg::PAR_BB) v = __ldg(p.dw_w + (i - Cfg::PAR_TAPS));
    else if (i < Cfg::PAR_GM) v = __ldg(p.dw_b + (i - Cfg::PAR_BB));
    else if (i < Cfg::PAR_WF) v = __ldg(p.gamma + (i - Cfg::PAR_GM));
    else if (i < Cfg::PAR_B2) v = __ldg(p.ffn_norm_w + (i - Cfg::PAR_WF));
    else if (i < Cfg::PAR_FG) v = __ldg(p.b2 + (i - Cfg::PAR_B2));
    else if (i < Cfg::PAR_B1) v = __ldg(p.ffn_gamma + (i - Cfg::PAR_FG));
    else v = __ldg(p.b1 + (i - Cfg::PAR_B1));
    par[i] = v;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_launch();

  if (warp == 0) {
    // ------------------------------ weight producer -----------------------------------
    // Weights never depend on a predecessor kernel: no griddepcontrol.wait here.
    if (lane == 0) {
      if constexpr (Cfg::RESIDENT) {
        mbar_arrive_expect_tx(w_full, Cfg::W_BYTES);
        for (int c = 0; c < NCH; ++c)
          tma_load_2d(wbuf + c * Cfg::W1_CHUNK, &tmW1, w_full, 0, c * HCH, VV_L2_EVICT_LAST);
        for (int c = 0; c < NCH; ++c)
          for (int kb = 0; kb < 2; ++kb)
            tma_load_2d(wbuf + NCH * Cfg::W1_CHUNK + c * Cfg::W2_CHUNK + kb * Cfg::W2_KB, &tmW2, w_full,
                        c * HCH + kb * 64, 0, VV_L2_EVICT_LAST);
      } else {
        // per tile, in the order the MMA warp consumes: W1(0), then for c: [W1(c+1)], W2(c); a unit = one 16 KB k-block
        int stage = 0;
        uint32_t phase = 0;
        auto push = [&](bool is_w1, int c) {
          for (int kb = 0; kb < 2; ++kb) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* dst = wbuf + stage * KB_BYTES;
            mbar_arrive_expect_tx(&full_bar[stage], KB_BYTES);
            if (is_w1) tma_load_2d(dst, &tmW1, &full_bar[stage], kb * 64, c * HCH, VV_L2_EVICT_LAST);
            else       tma_load_2d(dst, &tmW2, &full_bar[stage], c * HCH + kb * 64, 0, VV_L2_EVICT_LAST);
            if (++stage == NSTAGE) { stage = 0; phase ^= 1; }
          }
        };
        for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
          push(true, 0);
          for (int c = 0; c < NCH; ++c) {
            if (c + 1 < NCH) push(true, c + 1);
            push(false, c);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer ----------------------------------------
    // converged warp, elect.sync-issued (umma_*_warp, vv_common.cuh)
    constexpr uint32_t idesc1 = umma_idesc_bf16(BLK_M, HCH);
    constexpr uint32_t idesc2 = umma_idesc_bf16(BLK_M, C);
    const uint32_t y_addr0 = smem_u32(smem + Cfg::OFF_Y);
    const uint32_t h_addr0 = smem_u32(smem + Cfg::OFF_H);
    const uint32_t w_addr0 = smem_u32(wbuf);
    int stage = 0;
    uint32_t phase = 0;
    uint32_t it = 0, g = 0;
    if constexpr (Cfg::RESIDENT) {
      mbar_wait(w_full, 0);
      tc_fence_after();
    }
    auto mma1 = [&](int c, uint32_t gg, uint32_t y_addr) {   // D1[gg & 1] = Y . W1c^T
      const uint32_t d = tmem_base + (gg & 1) * HCH;
      if constexpr (Cfg::RESIDENT) {
        const uint32_t w_addr = w_addr0 + c * Cfg::W1_CHUNK;
#pragma unroll
        for (int kk = 0; kk < Cfg::NK1; ++kk)
          umma_bf16_warp(d, umma_desc_sw128(y_addr + kk * 32), umma_desc_sw128(w_addr + kk * 32), idesc1, kk != 0);
      } else {
#pragma unroll
        for (int kb = 0; kb < 2; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t w_addr = w_addr0 + stage * KB_BYTES;
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16_warp(d, umma_desc_sw128(y_addr + kb * KB_BYTES + kk * 32), umma_desc_sw128(w_addr + kk * 32),
                           idesc1, (kb | kk) != 0);
          umma_commit_warp(&empty_bar[stage]);
          if (++stage == NSTAGE) { stage = 0; phase ^= 1; }
        }
      }
      umma_commit_warp(&d1_full[gg & 1]);
    };
    auto mma2 = [&](int c, uint32_t gg, uint32_t tl) {   // D2[tl & 1] (+)= H . W2c^T
      const uint32_t hb = (Cfg::NH == 2) ? (gg & 1) : 0;
      const uint32_t h_addr = h_addr0 + hb * Cfg::H_BYTES;
      const uint32_t d = tmem_base + TMEM_D2 + (tl & 1) * C;
#pragma unroll
      for (int kb = 0; kb < 2; ++kb) {
        uint32_t w_addr;
        if constexpr (Cfg::RESIDENT) {
          w_addr = w_addr0 + NCH * Cfg::W1_CHUNK + c * Cfg::W2_CHUNK + kb * Cfg::W2_KB;
        } else {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          w_addr = w_addr0 + stage * KB_BYTES;
        }
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          umma_bf16_warp(d, umma_desc_sw128(h_addr + kb * KB_BYTES + kk * 32), umma_desc_sw128(w_addr + kk * 32), idesc2,
                         (c | kb | kk) != 0);
        if constexpr (!Cfg::RESIDENT) {
          umma_commit_warp(&empty_bar[stage]);
          if (++stage == NSTAGE) { stage = 0; phase ^= 1; }
        }
      }
      umma_commit_warp(&h_free[hb]);
      if (c == NCH - 1) umma_commit_warp(&d2_full[tl & 1]);
    };
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
      const uint32_t yb = it & 1;
      const uint32_t y_addr = y_addr0 + yb * Cfg::Y_BYTES;
      if (lane == 0 && it >= 1 && it <= 2) PTRACE(40 + (it - 1) * 4 + 0, clock64());
      mbar_wait_parked(&y_full[yb], (it >> 1) & 1);
      if (lane == 0 && it >= 1 && it <= 2) PTRACE(40 + (it - 1) * 4 + 1, clock64());
      tc_fence_after();
      mma1(0, g, y_addr);                  // D1[g&1] was drained before h_ready of chunk g-2 (waited on earlier)
      if (NCH == 1) umma_commit_warp(&y_free[yb]);
      for (int c = 0; c < NCH; ++c, ++g) {
        if (c + 1 < NCH) {
          mma1(c + 1, g + 1, y_addr);      // D1[(g+1)&1] was drained before h_ready of chunk g-1
          if (c + 1 == NCH - 1) umma_commit_warp(&y_free[yb]);
        }
        mbar_wait_parked(&h_ready[g & 1], (g >> 1) & 1);
        if (c == 0 && it >= 2) mbar_wait_parked(&d2_free[it & 1], ((it >> 1) + 1) & 1);   // epilogue 2 of tile it-2 has read D2[it & 1]
        if (lane == 0 && it >= 1 && it <= 2 && c == NCH - 1) PTRACE(40 + (it - 1) * 4 + 2, clock64());
        tc_fence_after();
        mma2(c, g, it);
        if (lane == 0 && it >= 1 && it <= 2 && c == NCH - 1) PTRACE(40 + (it - 1) * 4 + 3, clock64());
      }
    }
  } else if (warp < P_WARP_MIX0) {
    // ------------------------------ GELU / epilogue warps ------------------------------
    const int e = warp - P_WARP_EPI0;
    const int q = warp & 3;                        // TMEM lane quadrant of this warp
    const int qt = e >> 2;                         // which quarter of the columns (0..3)
    const int row = q * 32 + lane;                 // tile row this thread owns
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    griddep_wait();                                // x_out is the predecessor's input until now
    // ---- epilogue 2 of tile j: x'' = x' + ffn_gamma * (D2[j & 1] + b2), C/4 columns per thread.  It runs AFTER the
    // GELU chunks of tile j+1: D2 completes ~1 k clk after the tile's last H chunk, and with GELU work in between
    // nobody waits for it (before: 16 idle warps for 2 of 6 k clk per tile, r02 trace).
    constexpr int COLS2 = C / 4;
    const int col0 = qt * COLS2;
    auto epi2 = [&](int b, int t0, int rows_here, uint32_t j) {
      const bool quad_live = q * 32 < rows_here;
      const bool row_live = row < rows_here;
      bf16* orow = p.x_out + (static_cast<size_t>(b) * p.Tn + t0 + row) * C + col0;
      uint4 xv[COLS2 / 8];
#pragma unroll
      for (int k = 0; k < COLS2 / 8; ++k) xv[k] = row_live ? ldg_coherent(orow + k * 8) : make_uint4(0u, 0u, 0u, 0u);
      mbar_wait_parked(&d2_full[j & 1], (j >> 1) & 1);
      tc_fence_after();
      if (quad_live) {
        uint32_t r[COLS2];
        const uint32_t ta = lane_addr + TMEM_D2 + (j & 1) * C + col0;
        if constexpr (COLS2 == 8) tmem_ld_x8(ta, r);
        else if constexpr (COLS2 == 16) tmem_ld_x16(ta, r);
        else tmem_ld_32x32(ta, reinterpret_cast<uint32_t(&)[32]>(r));
        tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < COLS2 / 8; ++k) {
          float bv[8], gv[8], xr[8], v[8];
          lds8(par + Cfg::PAR_B2 + col0 + k * 8, bv);
          lds8(par + Cfg::PAR_FG + col0 + k * 8, gv);
          unpack_bf16x8(xv[k], xr);
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = fmaf(gv[i], __uint_as_float(r[k * 8 + i]) + bv[i], xr[i]);
          if (row_live) *reinterpret_cast<uint4*>(orow + k * 8) = pack_bf16x8(v);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&d2_free[j & 1]);
    };
    uint32_t it = 0, g = 0;
    int pb = 0, pt0 = 0, prows = 0;                // previous tile (its epilogue 2 is still owed)
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
      int b, t0;
      tile_coords(p, tile, b, t0);
      const int rows_here = min(BLK_M, p.Tn - t0);
      const bool quad_live = q * 32 < rows_here;   // this warp's 32 rows hold data (warp-uniform)
      const uint32_t yb = it & 1;
      // x' of this tile is in global memory once the mixer has signalled; tell the mixer this phase was seen
      mbar_wait_parked(&y_full[yb], (it >> 1) & 1);
      __syncwarp();
      if (lane == 0) mbar_arrive(&y_free[yb]);
      if (threadIdx.x == 64 && it < 5) PTRACE(it * 8 + 0, clock64());

      for (int c = 0; c < NCH; ++c, ++g) {
        mbar_wait_parked(&d1_full[g & 1], (g >> 1) & 1);
        if (threadIdx.x == 64 && it < 5 && c == 0) PTRACE(it * 8 + 1, clock64());
        tc_fence_after();
        uint4 pk[4];
        if (quad_live) {
          const float* b1 = par + Cfg::PAR_B1 + c * HCH + qt * 32;
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {         // 2 x 16 columns: keeps the accumulator words at 16 registers
            uint32_t r[16];
            tmem_ld_x16(lane_addr + (g & 1) * HCH + qt * 32 + hf * 16, r);
            tmem_ld_wait();
#pragma unroll
            for (int k = 0; k < 2; ++k) {          // chunks of 8 columns -> one 16-byte smem store
              float bv[8], v[8];
              lds8(b1 + hf * 16 + k * 8, bv);
              gelu8_bias(r + k * 8, bv, v);
              pk[hf * 2 + k] = pack_bf16x8(v);
            }
          }
        }
        // the MMA2 that last read this H buffer has finished
        uint32_t hb;
        if constexpr (Cfg::NH == 2) {
          hb = g & 1;
          if (g >= 2) mbar_wait_parked(&h_free[hb], ((g >> 1) + 1) & 1);
        } else {
          hb = 0;
          if (g >= 1) mbar_wait_parked(&h_free[0], (g - 1) & 1);
        }
        if (quad_live) {
          uint8_t* htile = smem + Cfg::OFF_H + hb * Cfg::H_BYTES + (qt >> 1) * KB_BYTES;   // columns 32*qt..: k-block qt/2
#pragma unroll
          for (int k = 0; k < 4; ++k) *reinterpret_cast<uint4*>(sw128(htile, row, (qt & 1) * 4 + k)) = pk[k];
        }
        tc_fence_before();                         // D1 reads ordered before the MMA warp reuses the buffer
        fence_proxy_async_smem();                  // generic-proxy smem writes -> visible to tcgen05.mma
        __syncwarp();
        if (lane == 0) mbar_arrive(&h_ready[g & 1]);
      }
      if (threadIdx.x == 64 && it < 5) PTRACE(it * 8 + 2, clock64());
      if (it > 0) epi2(pb, pt0, prows, it - 1);
      if (threadIdx.x == 64 && it < 5) PTRACE(it * 8 + 3, clock64());
      pb = b; pt0 = t0; prows = rows_here;
    }
    if (it > 0) epi2(pb, pt0, prows, it - 1);
  } else if (warp < P_WARP_ROWS) {
    // ------------------------------ mixer warps ----------------------------------------
    const int w = warp - P_WARP_MIX0;
    constexpr int LPR = C / 8;
    griddep_wait();                                // x_in / cache / x_out belong to the predecessor until now
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
      int b, t0;
      tile_coords(p, tile, b, t0);
      const int rows_here = min(BLK_M, p.Tn - t0);
      const uint32_t yb = it & 1, ph = (it >> 1) & 1;
      uint8_t* xb = smem + Cfg::OFF_XB + yb * Cfg::XB_BYTES;
      if (w == 0 && lane == 0 && it < 5) PTRACE(it * 8 + 4, clock64());
      mbar_wait_parked(&xb_full[yb], ph);
      if (w == 0 && lane == 0 && it < 5) PTRACE(it * 8 + 5, clock64());
      warp_mixer_norm<C>(p, smem, xb, t0, w, rows_here, lane);
      __syncwarp();
      if (lane == 0) mbar_arrive(&norm_done[yb * P_MIX_WARPS + w]);
      // ---- new history: normalised rows of time Tn-6 .. Tn-1 (the tile that holds the last row); a row is written
      // by the warp that normalised it (halo rows, when the utterance is shorter than 6 rows here: warp 0)
      if (t0 + BLK_M >= p.Tn) {
        bf16* dst = cache_wr(p.cv, b, p.cache_off);
        for (int i = lane; i < HIST * LPR; i += 32) {
          const int qh = i / LPR, l = i % LPR;
          const int r = p.Tn - t0 - HIST + qh;                 // tile row of time Tn-6+q (may be a halo row: < 0)
          const int owner = r < 0 ? 0 : r / WROWS;
          if (owner == w)
            *reinterpret_cast<uint4*>(dst + qh * C + l * 8) = *reinterpret_cast<const uint4*>(xb + (r + HIST) * (C * 2) + l * 16);
        }
      }
      // Y[yb] was last read by the MMA1s of tile it-2, and every epilogue warp has consumed that y_full phase
      if (it >= 2) mbar_wait_parked(&y_free[yb], ph ^ 1);
      if (w > 0) mbar_wait_parked(&norm_done[yb * P_MIX_WARPS + w - 1], ph);   // the 6 rows below this warp's first row
      if (w == 0 && lane == 0 && it < 5) PTRACE(it * 8 + 6, clock64());
      warp_mixer_conv<C>(p, smem, xb, smem + Cfg::OFF_Y + yb * Cfg::Y_BYTES, b, t0, w * WROWS, rows_here, lane);
      fence_proxy_async_smem();                    // generic-proxy smem writes -> visible to tcgen05.mma
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&y_full[yb]);
        mbar_arrive(&xb_free[yb]);
      }
      if (w == 0 && lane == 0 && it < 5) PTRACE(it * 8 + 7, clock64());
    }
  } else {
    // ------------------------------ row producer ---------------------------------------
    // one bulk copy (cp.async.bulk) per tile: its raw rows and the 6-row halo, two tiles ahead of the conv at most
    if (lane == 0) {
      griddep_wait();                              // x_in / cache belong to the predecessor until now
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
        int b, t0;
        tile_coords(p, tile, b, t0);
        const int rows_here = min(BLK_M, p.Tn - t0);
        const uint32_t yb = it & 1;
        uint8_t* xb = smem + Cfg::OFF_XB + yb * Cfg::XB_BYTES;
        if (it >= 2) mbar_wait_parked(&xb_free[yb], ((it >> 1) + 1) & 1);
        const uint32_t row_bytes = C * 2;
        mbar_arrive_expect_tx(&xb_full[yb], (rows_here + HIST) * row_bytes);
        const bf16* xu = p.x_in + static_cast<size_t>(b) * p.Tn * C;
        if (t0 == 0) {                             // rows before this call: already normalised, from the streaming cache
          bulk_g2s(xb, cache_rd(p.cv, b, p.cache_off), HIST * row_bytes, &xb_full[yb]);
          bulk_g2s(xb + HIST * row_bytes, xu, rows_here * row_bytes, &xb_full[yb]);
        } else {
          bulk_g2s(xb, xu + static_cast<size_t>(t0 - HIST) * C, (rows_here + HIST) * row_bytes, &xb_full[yb]);
        }
      }
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

template <int C>
int launch_pipe(const BlockFusedArgs& a, cudaStream_t stream) {
  using Cfg = PCfg<C>;
  static bool attr_set[64] = {};
  int dev = 0;
  VV_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) { vv_set_error("block_pipe: device ordinal %d", dev); return VV_ERR_UNSUPPORTED; }
  if (!attr_set[dev]) {           // function attributes are per device
    VV_CHECK_CUDA(cudaFuncSetAttribute(k_block_pipe<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    attr_set[dev] = true;
  }
  PipeParams p;
  p.x_in = a.x_in; p.x_out = a.x_out;
  p.norm_w = a.norm_w; p.dw_w = a.dw_w; p.dw_b = a.dw_b; p.gamma = a.gamma; p.ffn_norm_w = a.ffn_norm_w;
  p.b1 = a.b1; p.b2 = a.b2; p.ffn_gamma = a.ffn_gamma;
  p.cv = a.cv; p.cache_off = a.cache_off;
  p.n = a.n; p.Tn = a.Tn;
  p.full_per_utt = a.Tn / BLK_M;
  p.n_full = p.full_per_utt * a.n;
  p.n_tiles = p.n_full + ((a.Tn % BLK_M) ? a.n : 0);
  p.eps = a.eps;
  const int sms = vv_num_sms();
  const int grid = p.n_tiles < sms ? p.n_tiles : sms;
  VV_CHECK_CUDA(vv_launch(k_block_pipe<C>, dim3(grid), dim3(P_THREADS), Cfg::SMEM_BYTES, stream, 1, true,
                          *a.tm_w1, *a.tm_w2, p));
  return VV_OK;
}

}  // namespace

// Same contract as block_fused_launch (vv_block.cuh); the W2 tensor map's box is [C rows, 64 cols] in both.
int block_pipe_launch(const BlockFusedArgs& a, cudaStream_t stream) {
  VV_REQUIRE(a.x_in && a.x_out && a.x_in != a.x_out, "block_pipe: x_in/x_out must be distinct buffers");
  VV_REQUIRE(a.n >= 1 && a.Tn >= 1, "block_pipe: empty problem");
  switch (a.C) {
    case 32: return launch_pipe<32>(a, stream);
    case 64: return launch_pipe<64>(a, stream);
    case 128: return launch_pipe<128>(a, stream);
    default:
      vv_set_error("block_pipe: unsupported channel count %d", a.C);
      return VV_ERR_INVALID;
  }
}

int pipe_debug_trace(long long* out, int cap) {
#ifdef VV_GEMM_TRACE
  const int n = 160 * 48;
  if (cap < n) return -1;
  if (cudaMemcpyFromSymbol(out, g_pipe_trace, sizeof(long long) * n) != cudaSuccess) return -2;
  return n;
#else
  (void)out; (void)cap;
  return 0;
#endif
}
