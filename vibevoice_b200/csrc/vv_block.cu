// vv_block.cu — fused Block1D for the narrow decoder stages (C = 32 / 64 / 128).
//
// Reference: Block1D (modular_vibevoice_tokenizer.py:620-684, inlined in TokenizerDecoder
// .forward_features :924-944), ConvRMSNorm (:77-91), FFN (:579-596), SConv1d streaming (:327-382).
//
// Why: at n = 64 the last three stages hold 51 200 .. 204 800 time rows of 128 .. 32 channels.
// Unfused, every block writes and re-reads a 4C-wide hidden tensor and launches thousands of
// CTAs that each do one or two k-blocks of MMA (ncu r01: 125 .. 155 us per block).  Here one
// persistent CTA per SM walks 128-row tiles; per tile
//   compute warps : RMSNorm -> 7-tap depthwise conv (+6-row causal halo, from x or the cache)
//                   -> layer-scale residual x' (kept in smem) -> FFN RMSNorm -> Y, written straight
//                   into the UMMA K-major SWIZZLE_128B operand layout
//   MMA warp      : D1 = Y . W1c^T for each 128-wide chunk c of the hidden dim (TMEM, double buffered)
//   compute warps : tcgen05.ld D1 -> +b1 -> exact-erf GELU -> bf16 -> H in smem (operand layout)
//   MMA warp      : D2 += H . W2c^T  (TMEM)
//   compute warps : tcgen05.ld D2 -> +b2, *ffn_gamma, + x' -> bf16 -> global
// so the only HBM/L2 traffic is x in, x out and the weight chunks, which a TMA producer warp
// streams through a small ring (they are L2-resident: 16 KB .. 256 KB per block).
#include "vv_block.cuh"

#include "../../include/vvb200.h"
#include "vv_gemm.cuh"

namespace {

constexpr int BLK_M = 128;              // rows (time steps of one utterance) per tile
constexpr int HCH = 128;                // hidden-dim chunk
constexpr int N_CTRL_WARPS = 2;         // warp 0: TMA producer, warp 1: MMA issuer
constexpr int N_COMP_WARPS = 8;
constexpr int COMP_THREADS = 32 * N_COMP_WARPS;
constexpr int BLK_THREADS = 32 * (N_CTRL_WARPS + N_COMP_WARPS);
constexpr int KB_BYTES = BLK_M * 128;   // one [128 x 64] bf16 K-major SWIZZLE_128B tile

template <int C>
struct BCfg {
  static constexpr int NCH = 4 * C / HCH;                     // hidden chunks per tile: 1, 2, 4
  static constexpr int NKB1 = (C + 63) / 64;                  // 64-col k-blocks of Y / W1: 1, 1, 2
  static constexpr int NK1 = C / 16;                          // UMMA_K slices of MMA1
  static constexpr int W1_BYTES = NKB1 * KB_BYTES;            // [128 hidden x C(->64)]
  static constexpr int W2_BYTES = 2 * C * 128;                // 2 k-blocks of [C x 64]
  static constexpr int STAGE_BYTES = W1_BYTES > W2_BYTES ? W1_BYTES : W2_BYTES;
  static constexpr int NSTAGE = (C == 128) ? 3 : 4;
  static constexpr int XN_BYTES = (BLK_M + HIST) * C * 2;     // normalised rows incl. halo
  static constexpr int H_BYTES = 2 * KB_BYTES;                // [128 x 128] bf16
  static constexpr int XNH_BYTES = ((XN_BYTES > H_BYTES ? XN_BYTES : H_BYTES) + 1023) / 1024 * 1024;
  static constexpr int XP_PITCH = C * 2 + 16;                 // bytes; +16 keeps row-per-thread reads conflict-free
  static constexpr int XP_BYTES = BLK_M * XP_PITCH;           // x' (residual after the mixer)
  static constexpr int Y_BYTES = NKB1 * KB_BYTES;
  static constexpr int OFF_XNH = 0;                           // xn during the mixer phase, H afterwards
  static constexpr int OFF_Y = OFF_XNH + XNH_BYTES;
  static constexpr int OFF_RING = OFF_Y + Y_BYTES;
  static constexpr int OFF_XP = OFF_RING + NSTAGE * STAGE_BYTES;
  static constexpr int OFF_BAR = OFF_XP + XP_BYTES;
  static constexpr int SMEM_BYTES = 1024 + OFF_BAR + 256;
  static constexpr int TMEM_D1 = 0;                           // two [128 x 128] fp32 buffers
  static constexpr int TMEM_D2 = 2 * HCH;                     // [128 x C] fp32
  static constexpr int TMEM_COLS = 512;
};

#ifdef VV_GEMM_TRACE
__device__ long long g_ffn_trace[160][48];
#define FTRACE(slot, val) do { if ((slot) < 48) g_ffn_trace[blockIdx.x][slot] = (val); } while (0)
__device__ long long g_block_trace[160][48];
#define BTRACE(slot, val) do { if ((slot) < 48) g_block_trace[blockIdx.x][slot] = (val); } while (0)
#else
#define FTRACE(slot, val) do { } while (0)
#define BTRACE(slot, val) do { } while (0)
#endif

struct BlockKParams {
  const bf16* x_in; bf16* x_out;
  const float* norm_w; const float* dw_w; const float* dw_b; const float* gamma; const float* ffn_norm_w;
  const float* b1; const float* b2; const float* ffn_gamma;
  CacheView cv; size_t cache_off;
  int n, Tn, full_per_utt, n_full, n_tiles;   // tiles: the n_full whole 128-row tiles first, then one ragged tile per utterance
  float eps;
};

// 16 columns of 32 lanes (fp32): thread i of the warp receives lane (base+i).
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// Address of the 16-byte chunk (row, chunk) inside a [rows x 64] bf16 K-major SWIZZLE_128B tile
// (the layout TMA writes and tcgen05.mma reads: 128-byte rows, 16-byte chunks XORed with row & 7).
__device__ __forceinline__ uint8_t* sw128(uint8_t* tile, int row, int chunk) {
  return tile + row * 128 + ((chunk ^ (row & 7)) << 4);
}

__device__ __forceinline__ void named_bar_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}

// sum over the LPR lanes that share a row (LPR = C/8 <= 16: a sub-warp shuffle)
template <int LPR>
__device__ __forceinline__ float row_sum_lpr(float v) {
#pragma unroll
  for (int o = LPR >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void ld8(const float* p, float (&f)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}


// ---------------------------------------------------------------------------------------
// Mixer half of the block for one tile (rows t0 .. t0+rows_here-1 of utterance b), run by the
// 256 compute threads.  Thread layout as in k_block_dw (vv_dec.cu): 8 channels per thread,
// LPR = C/8 lanes per row, RPP = 256/LPR rows per pass.
//   xn  (smem, [(128+6) x C] bf16)  normalised rows incl. the 6-row causal halo
//   xp  (smem, [128] rows, pitch XP_PITCH)  x' = x + gamma * (dwconv(xn) + b)   (bf16: the stored stream)
//   Y   (smem, UMMA operand layout)  ConvRMSNorm_ffn(x')
// ---------------------------------------------------------------------------------------
template <int C>
__device__ __forceinline__ void mixer_phase(const BlockKParams& p, uint8_t* smem, int b, int t0, int rows_here,
                                            int tid) {
  using Cfg = BCfg<C>;
  constexpr int LPR = C / 8;
  constexpr int RPP = COMP_THREADS / LPR;
  bf16* xn = reinterpret_cast<bf16*>(smem + Cfg::OFF_XNH);
  uint8_t* xp = smem + Cfg::OFF_XP;
  uint8_t* ytile = smem + Cfg::OFF_Y;
  const int Tn = p.Tn;
  const int lane = tid % LPR;
  const int rslot = tid / LPR;
  const int c0 = lane * 8;
  const bf16* xin = p.x_in + static_cast<size_t>(b) * Tn * C;

  // ---- phase 1: normalised rows (time t0-6 .. t0+rows_here-1) into shared memory ----
  // All global loads of the tile are issued before the first use (one L2 round trip per tile
  // instead of one per pass); the raw rows are parked in the x' buffer for phase 2's residual.
  {
    constexpr int PASSES1 = (BLK_M + HIST + RPP - 1) / RPP;
    const int nrows = rows_here + HIST;
    uint4 raw[PASSES1];
#pragma unroll
    for (int ps = 0; ps < PASSES1; ++ps) {
      const int r = ps * RPP + rslot;
      const int tg = t0 - HIST + r;
      raw[ps] = make_uint4(0u, 0u, 0u, 0u);
      if (r < nrows) {
        if (tg >= 0) raw[ps] = *reinterpret_cast<const uint4*>(xin + static_cast<size_t>(tg) * C + c0);
        else         // rows before this call: already normalised, from the streaming cache
          raw[ps] = *reinterpret_cast<const uint4*>(cache_rd(p.cv, b, p.cache_off) + (HIST + tg) * C + c0);
      }
    }
    float wn[8];
    ld8(p.norm_w + c0, wn);
#pragma unroll
    for (int ps = 0; ps < PASSES1; ++ps) {
      const int r = ps * RPP + rslot;
      const int tg = t0 - HIST + r;
      const bool live = r < nrows;
      const bool from_x = live && tg >= 0;
      float f[8];
      float ss = 0.f;
      uint4 hv = raw[ps];
      if (from_x) {
        unpack_bf16x8(raw[ps], f);
#pragma unroll
        for (int i = 0; i < 8; ++i) ss += f[i] * f[i];
        if (r >= HIST) *reinterpret_cast<uint4*>(xp + (r - HIST) * Cfg::XP_PITCH + c0 * 2) = raw[ps];
      }
      ss = row_sum_lpr<LPR>(ss);
      if (from_x) {
        const float rms = rsqrtf(ss / static_cast<float>(C) + p.eps);
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = f[i] * rms * wn[i];
        hv = pack_bf16x8(f);
      }
      if (live) *reinterpret_cast<uint4*>(xn + static_cast<size_t>(r) * C + c0) = hv;
    }
  }
  named_bar_sync(1, COMP_THREADS);

  // ---- new history: normalised rows of time Tn-6 .. Tn-1 (the tile that holds the last row) ----
  if (t0 + BLK_M >= Tn) {
    bf16* dst = cache_wr(p.cv, b, p.cache_off);
    for (int i = tid; i < HIST * LPR; i += COMP_THREADS) {
      const int q = i / LPR, l = i % LPR;
      const int r = Tn - t0 + q;                           // smem row of time Tn-6+q
      *reinterpret_cast<uint4*>(dst + q * C + l * 8) =
          *reinterpret_cast<const uint4*>(xn + static_cast<size_t>(r) * C + l * 8);
    }
  }

  // ---- phase 2: 7-tap depthwise conv, layer-scale residual, FFN norm ----
  float wt[HIST + 1][8], bb[8], gm[8], wf[8];
#pragma unroll
  for (int j = 0; j <= HIST; ++j) ld8(p.dw_w + static_cast<size_t>(j) * C + c0, wt[j]);
  ld8(p.dw_b + c0, bb);
  ld8(p.gamma + c0, gm);
  ld8(p.ffn_norm_w + c0, wf);
  constexpr int PASSES2 = BLK_M / RPP;
#pragma unroll 2
  for (int ps = 0; ps < PASSES2; ++ps) {
    const int r = ps * RPP + rslot;
    const bool live = r < rows_here;
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
      unpack_bf16x8(*reinterpret_cast<const uint4*>(xp + r * Cfg::XP_PITCH + c0 * 2), xr);   // raw x[t] (phase 1)
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = xr[i] + gm[i] * acc[i];
      const uint4 ov = pack_bf16x8(o);
      *reinterpret_cast<uint4*>(xp + r * Cfg::XP_PITCH + c0 * 2) = ov;
      unpack_bf16x8(ov, o);                 // the FFN norm sees the stored (bf16) residual stream
#pragma unroll
      for (int i = 0; i < 8; ++i) ss += o[i] * o[i];
    }
    ss = row_sum_lpr<LPR>(ss);
    uint4 yv = make_uint4(0u, 0u, 0u, 0u);  // rows past the utterance end: zeros (their outputs are masked)
    if (live) {
      const float rms = rsqrtf(ss / static_cast<float>(C) + p.eps);
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = o[i] * rms * wf[i];
      yv = pack_bf16x8(o);
    }
    *reinterpret_cast<uint4*>(sw128(ytile + (c0 >> 6) * KB_BYTES, r, (c0 & 63) >> 3)) = yv;
  }
}


template <int C>
__global__ void __launch_bounds__(BLK_THREADS, 1)
k_block_fused(const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmW2,
              const BlockKParams p) {
  using Cfg = BCfg<C>;
  constexpr int NCH = Cfg::NCH;
  constexpr int NSTAGE = Cfg::NSTAGE;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* ring = smem + Cfg::OFF_RING;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::OFF_BAR);
  uint64_t* empty_bar = full_bar + NSTAGE;
  uint64_t* y_ready = empty_bar + NSTAGE;   // compute -> MMA: Y of this tile is in smem        (8 warps)
  uint64_t* d1_full = y_ready + 1;          // [2] MMA -> compute: D1 buffer holds chunk c       (commit)
  uint64_t* h_ready = d1_full + 2;          // compute -> MMA: H of chunk c in smem, D1 drained   (8 warps)
  uint64_t* h_free = h_ready + 1;           // MMA -> compute: MMA2 finished reading H            (commit)
  uint64_t* d2_full = h_free + 1;           // MMA -> compute: D2 of this tile complete           (commit)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(d2_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmW1);
    tma_prefetch_desc(&tmW2);
    for (int s = 0; s < NSTAGE; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(y_ready, N_COMP_WARPS);
    mbar_init(&d1_full[0], 1);
    mbar_init(&d1_full[1], 1);
    mbar_init(h_ready, N_COMP_WARPS);
    mbar_init(h_free, 1);
    mbar_init(d2_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_launch();

  if (warp == 0) {
    // ------------------------------ weight producer -----------------------------------
    // Per tile, in the order the MMA warp consumes: W1(0), then for c: [W1(c+1)], W2(c).
    // Weights never depend on a predecessor kernel: no griddepcontrol.wait here.
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      auto push = [&](bool is_w1, int c) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* dst = ring + stage * Cfg::STAGE_BYTES;
        if (is_w1) {
          mbar_arrive_expect_tx(&full_bar[stage], Cfg::W1_BYTES);
#pragma unroll
          for (int kb = 0; kb < Cfg::NKB1; ++kb)
            tma_load_2d(dst + kb * KB_BYTES, &tmW1, &full_bar[stage], kb * 64, c * HCH, VV_L2_EVICT_LAST);
        } else {
          mbar_arrive_expect_tx(&full_bar[stage], Cfg::W2_BYTES);
#pragma unroll
          for (int kb = 0; kb < 2; ++kb)
            tma_load_2d(dst + kb * (C * 128), &tmW2, &full_bar[stage], c * HCH + kb * 64, 0, VV_L2_EVICT_LAST);
        }
        if (++stage == NSTAGE) { stage = 0; phase ^= 1; }
      };
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        push(true, 0);
        for (int c = 0; c < NCH; ++c) {
          if (c + 1 < NCH) push(true, c + 1);
          push(false, c);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer ----------------------------------------
    // converged warp, elect.sync-issued (umma_*_warp, vv_common.cuh)
    {
      constexpr uint32_t idesc1 = umma_idesc_bf16(BLK_M, HCH);
      constexpr uint32_t idesc2 = umma_idesc_bf16(BLK_M, C);
      const uint32_t y_addr = smem_u32(smem + Cfg::OFF_Y);
      const uint32_t h_addr = smem_u32(smem + Cfg::OFF_XNH);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t it = 0, hcnt = 0;
      auto mma1 = [&](int c) {   // D1[c & 1] = Y . W1c^T
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t w_addr = smem_u32(ring + stage * Cfg::STAGE_BYTES);
        const uint32_t d = tmem_base + Cfg::TMEM_D1 + (c & 1) * HCH;
#pragma unroll
        for (int kk = 0; kk < Cfg::NK1; ++kk) {
          const uint32_t off = (kk >> 2) * KB_BYTES + (kk & 3) * 32;
          umma_bf16_warp(d, umma_desc_sw128(y_addr + off), umma_desc_sw128(w_addr + off), idesc1, kk != 0);
        }
        umma_commit_warp(&empty_bar[stage]);
        umma_commit_warp(&d1_full[c & 1]);
        if (++stage == NSTAGE) { stage = 0; phase ^= 1; }
      };
      auto mma2 = [&](int c) {   // D2 (+)= H . W2c^T
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t w_addr = smem_u32(ring + stage * Cfg::STAGE_BYTES);
        const uint32_t d = tmem_base + Cfg::TMEM_D2;
#pragma unroll
        for (int kk = 0; kk < HCH / 16; ++kk) {
          const uint32_t aoff = (kk >> 2) * KB_BYTES + (kk & 3) * 32;
          const uint32_t boff = (kk >> 2) * (C * 128) + (kk & 3) * 32;
          umma_bf16_warp(d, umma_desc_sw128(h_addr + aoff), umma_desc_sw128(w_addr + boff), idesc2, (c | kk) != 0);
        }
        umma_commit_warp(&empty_bar[stage]);
        umma_commit_warp(h_free);
        if (c == NCH - 1) umma_commit_warp(d2_full);
        if (++stage == NSTAGE) { stage = 0; phase ^= 1; }
      };
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
        mbar_wait(y_ready, it & 1);
        tc_fence_after();
        mma1(0);
        for (int c = 0; c < NCH; ++c) {
          if (c + 1 < NCH) mma1(c + 1);     // D1[(c+1)&1] was drained before h_ready of chunk c-1
          mbar_wait(h_ready, hcnt & 1);
          ++hcnt;
          tc_fence_after();
          mma2(c);
        }
      }
    }
  } else {
    // ------------------------------ compute warps -------------------------------------
    const int tid = threadIdx.x - 32 * N_CTRL_WARPS;
    const int q = warp & 3;                        // TMEM lane quadrant of this warp
    const int half = (warp - N_CTRL_WARPS) >> 2;   // which half of the columns
    const int row = q * 32 + lane;                 // tile row this thread owns in the epilogues
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    uint8_t* htile = smem + Cfg::OFF_XNH + half * KB_BYTES;   // H columns half*64 .. +63 = k-block `half`
    const uint8_t* xp_row = smem + Cfg::OFF_XP + row * Cfg::XP_PITCH;
    griddep_wait();                                // x_in / cache belong to the predecessor until now
    uint32_t it = 0, g = 0;                        // tiles done, hidden chunks done (global)
    uint32_t d1_uses[2] = {0u, 0u};
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
      // whole tiles are numbered first so that the short (ragged) tiles land on the CTAs that
      // got one whole tile fewer: balances the last wave
      int b, t0;
      if (tile < p.n_full) { b = tile / p.full_per_utt; t0 = (tile % p.full_per_utt) * BLK_M; }
      else                 { b = tile - p.n_full;       t0 = p.full_per_utt * BLK_M; }
      const int rows_here = min(BLK_M, p.Tn - t0);
      const bool quad_live = q * 32 < rows_here;   // this warp's 32 rows hold data (warp-uniform)
      if (tid == 0 && it < 5) BTRACE(it * 8 + 0, clock64());
      mixer_phase<C>(p, smem, b, t0, rows_here, tid);
      if (tid == 0 && it < 5) BTRACE(it * 8 + 1, clock64());
      fence_proxy_async_smem();                    // generic-proxy smem writes -> visible to tcgen05.mma
      __syncwarp();
      if (lane == 0) mbar_arrive(y_ready);

      for (int c = 0; c < NCH; ++c, ++g) {
        mbar_wait(&d1_full[c & 1], d1_uses[c & 1] & 1);
        if (tid == 0 && it < 5 && c == 0) BTRACE(it * 8 + 2, clock64());
        ++d1_uses[c & 1];
        if (g > 0) mbar_wait(h_free, (g - 1) & 1);  // MMA2 of the previous chunk has finished reading H
        tc_fence_after();
        const float* b1 = p.b1 + c * HCH + half * 64;
#pragma unroll
        for (int j = 0; j < (quad_live ? 2 : 0); ++j) {   // 2 x 32 columns of D1
          uint32_t r[32];
          tmem_ld_32x32(lane_addr + Cfg::TMEM_D1 + (c & 1) * HCH + half * 64 + j * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int k = 0; k < 4; ++k) {            // chunks of 8 columns -> one 16-byte smem store
            float bv[8], v[8];
            ld8(b1 + j * 32 + k * 8, bv);
            gelu8_bias(r + k * 8, bv, v);
            *reinterpret_cast<uint4*>(sw128(htile, row, j * 4 + k)) = pack_bf16x8(v);
          }
        }
        tc_fence_before();                         // D1 reads ordered before the MMA warp reuses it
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(h_ready);
        if (tid == 0 && it < 5 && c == NCH - 1) BTRACE(it * 8 + 3, clock64());
      }

      // ---- epilogue 2: x'' = x' + ffn_gamma * (D2 + b2) ----
      mbar_wait(d2_full, it & 1);
      if (tid == 0 && it < 5) BTRACE(it * 8 + 4, clock64());
      tc_fence_after();
      constexpr int COLS2 = C / 2;                 // columns per thread (two warps share a quadrant)
      const int col0 = half * COLS2;
      bf16* orow = p.x_out + (static_cast<size_t>(b) * p.Tn + t0 + row) * C + col0;
#pragma unroll
      for (int j = 0; j < (quad_live ? COLS2 / 16 : 0); ++j) {
        uint32_t r[16];
        tmem_ld_32x16(lane_addr + Cfg::TMEM_D2 + col0 + j * 16, r);
        tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          const int col = col0 + j * 16 + k * 8;
          float bv[8], gv[8], xr[8], v[8];
          ld8(p.b2 + col, bv);
          ld8(p.ffn_gamma + col, gv);
          unpack_bf16x8(*reinterpret_cast<const uint4*>(xp_row + col * 2), xr);
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = xr[i] + gv[i] * (__uint_as_float(r[k * 8 + i]) + bv[i]);
          if (row < rows_here) *reinterpret_cast<uint4*>(orow + j * 16 + k * 8) = pack_bf16x8(v);
        }
      }
      tc_fence_before();
      // next tile's mixer overwrites xn/H, Y and x': all MMAs of this tile are complete (d2_full),
      // but every compute warp must be done reading x' and D2 first.
      named_bar_sync(2, COMP_THREADS);
      if (tid == 0 && it < 5) BTRACE(it * 8 + 5, clock64());
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}


// ---------------------------------------------------------------------------------------
// FFN half of a Block1D for C = 256 (stage 3: 12 800 rows at n = 64), fused so that the 4C-wide hidden
// tensor never leaves the SM:      x'' = x' + ffn_gamma * (W2 . gelu(W1 . y + b1) + b2)
// y = ConvRMSNorm_ffn(x') and x' come from k_block_dw (vv_dec.cu).  The mixer does not fit next to a
// [128 x 256] operand tile plus its halo, so this kernel takes Y by TMA instead of computing it.
// Unfused, the two GEMMs moved 2 x 26 MB of hidden activations per block through L2/HBM and ran
// at ~130 TFLOP/s (51 + 26 us per block, ncu r01); here a tile's traffic is y in, x' in, x'' out and
// 1 MB of L2-resident weights.
//   warp 0  : TMA producer - Y tile (4 boxes, own barrier), then W1/W2 chunks through a 4 x 32 KB ring
//   warp 1  : MMA issuer   - D1[c&1] = Y . W1c^T (N = 128, K = 256) ; D2 += H . W2c^T (N = 256, K = 128)
//   warps 2-9: epilogue 1 (tcgen05.ld D1 -> +b1 -> GELU -> H in the UMMA operand layout) and
//              epilogue 2 (tcgen05.ld D2 -> +b2, *gamma, + x' -> global)
// TMEM: D1 2 x 128 columns, D2 256 columns (all 512).  The next tile's Y load and first MMA1s overlap
// this tile's last MMA2s and epilogue 2.
// ---------------------------------------------------------------------------------------
constexpr int FC = 256;
constexpr int F_NCH = 4 * FC / HCH;            // 8 hidden chunks
constexpr int F_NKB = FC / 64;                 // 4 k-blocks of Y / W1
constexpr int F_STAGE_BYTES = 32 * 1024;       // W1: 2 k-blocks [128 x 64]; W2: 1 k-block [256 x 64]
constexpr int F_NSTAGE = 4;
constexpr int F_COMP_WARPS = 16;             // 4 per scheduler: the GELU epilogue is latency-bound with 2
constexpr int F_THREADS = 32 * (N_CTRL_WARPS + F_COMP_WARPS);
constexpr int F_OFF_Y = 0;
constexpr int F_OFF_H = F_OFF_Y + F_NKB * KB_BYTES;
constexpr int F_OFF_RING = F_OFF_H + 2 * KB_BYTES;
constexpr int F_OFF_BAR = F_OFF_RING + F_NSTAGE * F_STAGE_BYTES;
constexpr int F_SMEM_BYTES = 1024 + F_OFF_BAR + 256;
static_assert(F_SMEM_BYTES <= 227 * 1024, "k_ffn_fused shared memory");

struct FfnKParams {
  const bf16* xres; bf16* x_out;               // x' and x'' ([rows, 256] bf16; may alias: same thread reads then writes)
  const float* b1; const float* b2; const float* ffn_gamma;
  int rows, n_tiles;
};

__global__ void __launch_bounds__(F_THREADS, 1)
k_ffn_fused(const __grid_constant__ CUtensorMap tmY, const __grid_constant__ CUtensorMap tmW1,
            const __grid_constant__ CUtensorMap tmW2, const FfnKParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* ring = smem + F_OFF_RING;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + F_OFF_BAR);
  uint64_t* empty_bar = full_bar + F_NSTAGE;
  uint64_t* y_full = empty_bar + F_NSTAGE;  // TMA -> MMA: Y tile landed
  uint64_t* y_free = y_full + 1;            // MMA -> TMA: last MMA1 of the tile has read Y       (commit)
  uint64_t* d1_full = y_free + 1;           // [2] MMA -> compute                                 (commit)
  uint64_t* h_ready = d1_full + 2;          // compute -> MMA: H of chunk c in smem, D1 drained   (8 warps)
  uint64_t* h_free = h_ready + 1;           // MMA -> compute: MMA2 finished reading H            (commit)
  uint64_t* d2_full = h_free + 1;           // MMA -> compute: D2 of this tile complete           (commit)
  uint64_t* d2_free = d2_full + 1;          // compute -> MMA: D2 of this tile has been read      (8 warps)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(d2_free + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmY);
    tma_prefetch_desc(&tmW1);
    tma_prefetch_desc(&tmW2);
    for (int s = 0; s < F_NSTAGE; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(y_full, 1);
    mbar_init(y_free, 1);
    mbar_init(&d1_full[0], 1);
    mbar_init(&d1_full[1], 1);
    mbar_init(h_ready, F_COMP_WARPS);
    mbar_init(h_free, 1);
    mbar_init(d2_full, 1);
    mbar_init(d2_free, F_COMP_WARPS);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  constexpr uint32_t T_D1 = 0, T_D2 = 2 * HCH;
  griddep_launch();

  if (warp == 0) {
    // ------------------------------ TMA producer --------------------------------------
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      auto push_w1 = [&](int c) {            // two stages: k-blocks {0,1} and {2,3} of W1 rows c*128..
        for (int s2 = 0; s2 < 2; ++s2) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* dst = ring + stage * F_STAGE_BYTES;
          mbar_arrive_expect_tx(&full_bar[stage], F_STAGE_BYTES);
          tma_load_2d(dst, &tmW1, &full_bar[stage], (2 * s2) * 64, c * HCH, VV_L2_EVICT_LAST);
          tma_load_2d(dst + KB_BYTES, &tmW1, &full_bar[stage], (2 * s2 + 1) * 64, c * HCH, VV_L2_EVICT_LAST);
          if (++stage == F_NSTAGE) { stage = 0; phase ^= 1; }
        }
      };
      auto push_w2 = [&](int c) {            // two stages: hidden columns c*128 + {0..63} and {64..127}, all 256 rows
        for (int s2 = 0; s2 < 2; ++s2) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], F_STAGE_BYTES);
          tma_load_2d(ring + stage * F_STAGE_BYTES, &tmW2, &full_bar[stage], c * HCH + s2 * 64, 0, VV_L2_EVICT_LAST);
          if (++stage == F_NSTAGE) { stage = 0; phase ^= 1; }
        }
      };
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
        push_w1(0);                          // weights never depend on the predecessor kernel
        if (it == 0) griddep_wait();         // Y does
        else mbar_wait(y_free, (it - 1) & 1);
        mbar_arrive_expect_tx(y_full, F_NKB * KB_BYTES);
        for (int kb = 0; kb < F_NKB; ++kb)
          tma_load_2d(smem + F_OFF_Y + kb * KB_BYTES, &tmY, y_full, kb * 64, tile * BLK_M, VV_L2_EVICT_FIRST);
        for (int c = 0; c < F_NCH; ++c) {
          if (c + 1 < F_NCH) push_w1(c + 1);
          push_w2(c);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer ----------------------------------------
    // converged warp, elect.sync-issued (umma_*_warp, vv_common.cuh)
    {
      constexpr uint32_t idesc1 = umma_idesc_bf16(BLK_M, HCH);
      constexpr uint32_t idesc2 = umma_idesc_bf16(BLK_M, FC);
      const uint32_t y_addr = smem_u32(smem + F_OFF_Y);
      const uint32_t h_addr = smem_u32(smem + F_OFF_H);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t it = 0, hcnt = 0;
      auto mma1 = [&](int c, bool last) {    // D1[c & 1] = Y . W1c^T
        const uint32_t d = tmem_base + T_D1 + (c & 1) * HCH;
        for (int s2 = 0; s2 < 2; ++s2) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t w_addr = smem_u32(ring + stage * F_STAGE_BYTES);
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint32_t aoff = (2 * s2 + (kk >> 2)) * KB_BYTES + (kk & 3) * 32;
            const uint32_t boff = (kk >> 2) * KB_BYTES + (kk & 3) * 32;
            umma_bf16_warp(d, umma_desc_sw128(y_addr + aoff), umma_desc_sw128(w_addr + boff), idesc1, (s2 | kk) != 0);
          }
          umma_commit_warp(&empty_bar[stage]);
          if (++stage == F_NSTAGE) { stage = 0; phase ^= 1; }
        }
        umma_commit_warp(&d1_full[c & 1]);
        if (last) umma_commit_warp(y_free);
      };
      auto mma2 = [&](int c) {               // D2 (+)= H . W2c^T
        const uint32_t d = tmem_base + T_D2;
        for (int s2 = 0; s2 < 2; ++s2) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t w_addr = smem_u32(ring + stage * F_STAGE_BYTES);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16_warp(d, umma_desc_sw128(h_addr + s2 * KB_BYTES + kk * 32), umma_desc_sw128(w_addr + kk * 32), idesc2,
                      (c | s2 | kk) != 0);
          umma_commit_warp(&empty_bar[stage]);
          if (++stage == F_NSTAGE) { stage = 0; phase ^= 1; }
        }
        umma_commit_warp(h_free);
        if (c == F_NCH - 1) umma_commit_warp(d2_full);
      };
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
        mbar_wait(y_full, it & 1);
        tc_fence_after();
        mma1(0, false);
        for (int c = 0; c < F_NCH; ++c) {
          if (c + 1 < F_NCH) mma1(c + 1, c + 1 == F_NCH - 1);   // D1[(c+1)&1] was drained before h_ready of chunk c-1
          mbar_wait(h_ready, hcnt & 1);
          ++hcnt;
          if (c == 0 && it > 0) mbar_wait(d2_free, (it - 1) & 1);   // epilogue 2 of the previous tile has read D2
          tc_fence_after();
          mma2(c);
        }
      }
    }
  } else {
    // ------------------------------ compute warps -------------------------------------
    const int q = warp & 3;                        // TMEM lane quadrant of this warp
    const int qt = (warp - N_CTRL_WARPS) >> 2;     // which quarter of the columns (0..3)
    const int row = q * 32 + lane;
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    uint8_t* htile = smem + F_OFF_H + (qt >> 1) * KB_BYTES;   // H columns 32*qt .. +31: k-block qt/2, chunks 4*(qt&1)..
    griddep_wait();                                // x' belongs to the predecessor until now
    uint32_t it = 0, g = 0;
    uint32_t d1_uses[2] = {0u, 0u};
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
      const int grow = tile * BLK_M + row;
      const bool quad_live = tile * BLK_M + q * 32 < p.rows;   // warp-uniform
      for (int c = 0; c < F_NCH; ++c, ++g) {
        if (threadIdx.x == 64 && it == 0 && c == 0) FTRACE(46, clock64());
        mbar_wait(&d1_full[c & 1], d1_uses[c & 1] & 1);
        ++d1_uses[c & 1];
        if (threadIdx.x == 64 && it == 0) FTRACE(c * 4, clock64());
        if (g > 0) mbar_wait(h_free, (g - 1) & 1);
        if (threadIdx.x == 64 && it == 0) FTRACE(c * 4 + 1, clock64());
        tc_fence_after();
        if (quad_live) {
          const float* b1 = p.b1 + c * HCH + qt * 32;
          uint32_t r[32];
          tmem_ld_32x32(lane_addr + T_D1 + (c & 1) * HCH + qt * 32, r);
          tmem_ld_wait();
          if (threadIdx.x == 64 && it == 0 && c >= 4) FTRACE(32 + (c - 4) * 2, clock64());
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            float bv[8], v[8];
            ld8(b1 + k * 8, bv);
            gelu8_bias(r + k * 8, bv, v);
            *reinterpret_cast<uint4*>(sw128(htile, row, (qt & 1) * 4 + k)) = pack_bf16x8(v);
          }
          if (threadIdx.x == 64 && it == 0 && c >= 4) FTRACE(33 + (c - 4) * 2, clock64());
        }
        tc_fence_before();
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(h_ready);
        if (threadIdx.x == 64 && it == 0) FTRACE(c * 4 + 2, clock64());
      }
      // ---- epilogue 2: x'' = x' + ffn_gamma * (D2 + b2), 64 columns per thread ----
      mbar_wait(d2_full, it & 1);
      if (threadIdx.x == 64 && it == 0) FTRACE(40, clock64());
      tc_fence_after();
      const int col0 = qt * (FC / 4);
      const bf16* xrow = p.xres + static_cast<size_t>(grow) * FC + col0;
      bf16* orow = p.x_out + static_cast<size_t>(grow) * FC + col0;
      const bool row_live = grow < p.rows;
#pragma unroll 2
      for (int j = 0; j < (quad_live ? FC / 4 / 16 : 0); ++j) {
        uint4 xv[2];
        xv[0] = row_live ? *reinterpret_cast<const uint4*>(xrow + j * 16) : make_uint4(0u, 0u, 0u, 0u);
        xv[1] = row_live ? *reinterpret_cast<const uint4*>(xrow + j * 16 + 8) : make_uint4(0u, 0u, 0u, 0u);
        uint32_t r[16];
        tmem_ld_32x16(lane_addr + T_D2 + col0 + j * 16, r);
        tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          const int col = col0 + j * 16 + k * 8;
          float bv[8], gv[8], xr[8], v[8];
          ld8(p.b2 + col, bv);
          ld8(p.ffn_gamma + col, gv);
          unpack_bf16x8(xv[k], xr);
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = xr[i] + gv[i] * (__uint_as_float(r[k * 8 + i]) + bv[i]);
          if (row_live) *reinterpret_cast<uint4*>(orow + j * 16 + k * 8) = pack_bf16x8(v);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(d2_free);
      if (threadIdx.x == 64 && it == 0) FTRACE(41, clock64());
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

bool block_fused_supported(int C, int Tn) { return (C == 32 || C == 64 || C == 128) && Tn >= 1; }

int block_fused_make_tmaps(const bf16* w1, const bf16* w2, int C, CUtensorMap* tm_w1, CUtensorMap* tm_w2) {
  VV_TRY(gemm_make_tmap(tm_w1, w1, 4 * C, C, C, HCH));        // W1 [4C, C]: boxes of 128 hidden rows
  VV_TRY(gemm_make_tmap(tm_w2, w2, C, 4 * C, 4 * C, C));      // W2 [C, 4C]: boxes of all C rows
  return VV_OK;
}

namespace {
template <int C>
int launch_c(const BlockFusedArgs& a, cudaStream_t stream) {
  using Cfg = BCfg<C>;
  static bool attr_flags[VV_MAX_DEVICES] = {};
  bool& attr_set = attr_flags[vv_current_device()];   // function attributes are per device
  if (!attr_set) {
    VV_CHECK_CUDA(cudaFuncSetAttribute(k_block_fused<C>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg::SMEM_BYTES));
    attr_set = true;
  }
  BlockKParams p;
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
  VV_CHECK_CUDA(vv_launch(k_block_fused<C>, dim3(grid), dim3(BLK_THREADS), Cfg::SMEM_BYTES, stream, 1, true,
                          *a.tm_w1, *a.tm_w2, p));
  return VV_OK;
}
}  // namespace

int block_fused_launch(const BlockFusedArgs& a, cudaStream_t stream) {
  VV_REQUIRE(a.x_in && a.x_out && a.x_in != a.x_out, "block_fused: x_in/x_out must be distinct buffers");
  VV_REQUIRE(a.n >= 1 && a.Tn >= 1, "block_fused: empty problem");
  if (vv_option_block_pipe()) return block_pipe_launch(a, stream);
  switch (a.C) {
    case 32: return launch_c<32>(a, stream);
    case 64: return launch_c<64>(a, stream);
    case 128: return launch_c<128>(a, stream);
    default:
      vv_set_error("block_fused: unsupported channel count %d", a.C);
      return VV_ERR_INVALID;
  }
}

int block_debug_trace(long long* out, int cap) {
#ifdef VV_GEMM_TRACE
  const int n = 160 * 48;
  if (cap < n) return -1;
  if (cudaMemcpyFromSymbol(out, g_block_trace, sizeof(long long) * n) != cudaSuccess) return -2;
  return n;
#else
  (void)out; (void)cap;
  return 0;
#endif
}

int ffn_debug_trace(long long* out, int cap) {
#ifdef VV_GEMM_TRACE
  const int n = 160 * 48;
  if (cap < n) return -1;
  if (cudaMemcpyFromSymbol(out, g_ffn_trace, sizeof(long long) * n) != cudaSuccess) return -2;
  return n;
#else
  (void)out; (void)cap;
  return 0;
#endif
}

// ---- fused FFN (C = 256) ----
bool ffn_fused_supported(int C) { return C == FC; }

int ffn_fused_make_tmaps(const bf16* w1, const bf16* w2, int C, CUtensorMap* tm_w1, CUtensorMap* tm_w2) {
  VV_REQUIRE(C == FC, "ffn_fused: unsupported channel count %d", C);
  VV_TRY(gemm_make_tmap(tm_w1, w1, 4 * C, C, C, HCH));        // W1 [4C, C]: boxes of 128 hidden rows x 64 cols
  VV_TRY(gemm_make_tmap(tm_w2, w2, C, 4 * C, 4 * C, C));      // W2 [C, 4C]: boxes of all 256 rows x 64 cols
  return VV_OK;
}

int ffn_fused_launch(const FfnFusedArgs& a, cudaStream_t stream) {
  VV_REQUIRE(a.y && a.xres && a.x_out && a.rows >= 1 && a.C == FC, "ffn_fused: bad arguments");
  static bool attr_flags[VV_MAX_DEVICES] = {};
  bool& attr_set = attr_flags[vv_current_device()];   // function attributes are per device
  if (!attr_set) {
    VV_CHECK_CUDA(cudaFuncSetAttribute(k_ffn_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, F_SMEM_BYTES));
    attr_set = true;
  }
  CUtensorMap tmY;
  VV_TRY(gemm_make_tmap(&tmY, a.y, a.rows, FC, FC, BLK_M));   // rows past the end are zero-filled by TMA
  FfnKParams p;
  p.xres = a.xres; p.x_out = a.x_out; p.b1 = a.b1; p.b2 = a.b2; p.ffn_gamma = a.ffn_gamma;
  p.rows = a.rows;
  p.n_tiles = vv_ceil_div(a.rows, BLK_M);
  const int sms = vv_num_sms();
  const int grid = p.n_tiles < sms ? p.n_tiles : sms;
  VV_CHECK_CUDA(vv_launch(k_ffn_fused, dim3(grid), dim3(F_THREADS), F_SMEM_BYTES, stream, 1, true, tmY, *a.tm_w1,
                          *a.tm_w2, p));
  return VV_OK;
}
