// vv_gemm.cu — tcgen05 / TMEM / TMA GEMM with fused epilogues (see vv_gemm.cuh).
#include "vv_gemm.cuh"
#include "vv_gemm_dev.cuh"

#include <mutex>
#include <vector>

namespace {

#ifdef VV_GEMM_TRACE
// Debug build only (make TRACE=1): per-CTA clock64 stamps, read back with vv_debug_trace().
__device__ long long g_gemm_trace[160][48];
#define TRACE(slot, val) do { if (blockIdx.x < 160 && blockIdx.y == 0 && blockIdx.z == 0 && (slot) < 48) g_gemm_trace[blockIdx.x][slot] = (val); } while (0)
#else
#define TRACE(slot, val) do { } while (0)
#endif

template <int BN>
struct TileCfg {
  static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;
  static constexpr int B_BYTES = BN * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int TMEM_COLS = BN < 32 ? 32 : BN;
  static constexpr int MAX_STAGES = 12;
  static constexpr int smem_bytes(int stages) { return 1024 + stages * STAGE_BYTES + 256; }
};

// CL = cluster size along N (1, 2, 4).  The CL CTAs of a cluster compute adjacent N tiles of the
// same M tile: each loads 128/CL rows of the A (activation) tile and multicasts them to all, so
// the L2 -> SM traffic for A drops by CL.  That matters for the weight-streaming launches
// (M <= 128): without it every CTA re-reads the whole A tile for a few KB of weights and the
// kernel is bound by L2 output bandwidth instead of HBM.
//
// PDL: B is a weight matrix, never written by a kernel, so the producer fills the first
// `stages` B slots BEFORE griddepcontrol.wait; A loads and everything that touches
// activations come after it.  With the predecessor still draining, the HBM stream of this
// launch is already running.
//
// SWAP (skinny problems, M <= BN rows): the roles of the operands are exchanged - the WEIGHT matrix feeds the
// 128-row M side of the MMA (tmA = weights, boxes of 128 output features) and the activations the BN-wide N side
// (tmB = activations, boxes of BN rows), so D holds out^T: TMEM lane = output feature, column = batch row.  With
// M <= 64 rows a plain tile is half padding; swapped, every MMA instruction and every byte of the ring is useful,
// which halves the k-loop of an issue-bound launch (profiles/r01_mma_issue.md).  blockIdx.x = feature tile,
// blockIdx.y = row tile; the epilogue stores one batch row per instruction (32 lanes = 32 consecutive features).
template <int BN, int CL, bool SWAP = false>
__global__ void __launch_bounds__(NUM_THREADS, 2)
vv_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const GemmKParams p) {
  static_assert(!SWAP || CL == 1, "swapped operands: no multicast");
  using Cfg = TileCfg<BN>;
  const int STAGES = p.stages;
  constexpr int A_SLICE_ROWS = BLOCK_M / CL;
  constexpr int A_SLICE_BYTES = Cfg::A_BYTES / CL;
  constexpr uint16_t CL_MASK = static_cast<uint16_t>((1u << CL) - 1);
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B operands need 1024-byte aligned tiles (dynamic smem starts at the same
  // CTA-relative offset in every CTA of the cluster, so multicast offsets agree).
  uint8_t* smem = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * Cfg::A_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* accum_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n_blk = blockIdx.x, m_blk = blockIdx.y, z = blockIdx.z;
  const int crank = (CL > 1) ? static_cast<int>(cluster_ctarank()) : 0;

  const int kb_total = (p.K + BLOCK_K - 1) / BLOCK_K;
  const int kb0 = z * p.kb_per_split;
  const int kb1 = min(kb_total, kb0 + p.kb_per_split);
  const int nkb = max(kb1 - kb0, 0);
  if (threadIdx.x == 0) TRACE(0, clock64());

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], CL);     // one tcgen05.commit per CTA of the cluster
    }
    mbar_init(accum_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  tc_fence_before();
  if (CL > 1) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_launch();                     // let the next kernel in the stream start its prologue
  if (threadIdx.x == 0) TRACE(1, clock64());

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (lane == 0) {
      const int npre = min(STAGES, nkb);
      const uint32_t stage_tx = (CL > 1 || SWAP) ? Cfg::STAGE_BYTES : static_cast<uint32_t>(p.a_rows * BLOCK_K * 2 + Cfg::B_BYTES);
      for (int i = 0; i < npre; ++i) {        // weights first: independent of the predecessor
        mbar_arrive_expect_tx(&full_bar[i], stage_tx);
        if (SWAP)
          tma_load_2d(smem_a + i * Cfg::A_BYTES, &tmA, &full_bar[i], (kb0 + i) * BLOCK_K, n_blk * BLOCK_M, p.b_policy);
        else
          tma_load_2d(smem_b + i * Cfg::B_BYTES, &tmB, &full_bar[i], (kb0 + i) * BLOCK_K, n_blk * BN,
                      p.b_policy);
      }
      TRACE(2, clock64());
      griddep_wait();
      TRACE(3, clock64());
      for (int i = 0; i < npre; ++i) {
        uint8_t* dst = smem_a + i * Cfg::A_BYTES + crank * A_SLICE_BYTES;
        if (SWAP)
          tma_load_2d(smem_b + i * Cfg::B_BYTES, &tmB, &full_bar[i], (kb0 + i) * BLOCK_K, m_blk * BN, p.a_policy);
        else if (CL > 1)
          tma_load_2d_mc(dst, &tmA, &full_bar[i], (kb0 + i) * BLOCK_K,
                         m_blk * BLOCK_M + crank * A_SLICE_ROWS, CL_MASK, p.a_policy);
        else
          tma_load_2d(dst, &tmA, &full_bar[i], (kb0 + i) * BLOCK_K, m_blk * BLOCK_M, p.a_policy);
      }
      int stage = (npre == STAGES) ? 0 : npre;
      uint32_t phase = (npre == STAGES) ? 1 : 0;
      for (int kb = kb0 + npre; kb < kb1; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);   // every CTA of the cluster has drained the slot
        mbar_arrive_expect_tx(&full_bar[stage], stage_tx);
        uint8_t* dst = smem_a + stage * Cfg::A_BYTES + crank * A_SLICE_BYTES;
        if (SWAP) {
          tma_load_2d(smem_a + stage * Cfg::A_BYTES, &tmA, &full_bar[stage], kb * BLOCK_K, n_blk * BLOCK_M, p.b_policy);
          tma_load_2d(smem_b + stage * Cfg::B_BYTES, &tmB, &full_bar[stage], kb * BLOCK_K, m_blk * BN, p.a_policy);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
          continue;
        }
        if (CL > 1)
          tma_load_2d_mc(dst, &tmA, &full_bar[stage], kb * BLOCK_K,
                         m_blk * BLOCK_M + crank * A_SLICE_ROWS, CL_MASK, p.a_policy);
        else
          tma_load_2d(dst, &tmA, &full_bar[stage], kb * BLOCK_K, m_blk * BLOCK_M, p.a_policy);
        tma_load_2d(smem_b + stage * Cfg::B_BYTES, &tmB, &full_bar[stage], kb * BLOCK_K,
                    n_blk * BN, p.b_policy);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer --------------------------------
    // The whole warp runs the loop converged and elect.sync picks the issuing lane (umma_*_warp): issued
    // from `if (lane == 0)` every UTCHMMA / UTCBAR sits in an ELECT..BRA.U.ANY loop (profiles/r01_mma_issue.md).
    {
      constexpr uint32_t idesc = umma_idesc_bf16(BLOCK_M, BN);
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        if (lane == 0) TRACE(8 + kb, clock64());
        tc_fence_after();
        const uint64_t adesc = umma_desc_sw128(smem_u32(smem_a + stage * Cfg::A_BYTES));
        const uint64_t bdesc = umma_desc_sw128(smem_u32(smem_b + stage * Cfg::B_BYTES));
#pragma unroll
        for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
          // advance 32 B (= UMMA_K bf16) inside the 128-B swizzle row: +2 in the >>4 field
          umma_bf16_warp(tmem_base, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
        }
        // frees the smem slot (in every CTA of the cluster) when these MMAs retire
        if (CL > 1) umma_commit_mc_warp(&empty_bar[stage], CL_MASK);
        else umma_commit_warp(&empty_bar[stage]);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      umma_commit_warp(accum_bar);        // accumulator complete
    }
  } else {
    // ------------------------------ epilogue ----------------------------------
    const int q = warp & 3;               // TMEM lane quadrant this warp may read
    const int half = (warp - 2) >> 2;     // two warps share a quadrant; split the columns
    griddep_wait();                       // resid / gate / out belong to the predecessor until now
    mbar_wait(accum_bar, 0);
    if (threadIdx.x == 64) TRACE(4, clock64());
    tc_fence_after();
    const int row = m_blk * BLOCK_M + q * 32 + lane;
    constexpr int NCHUNK = (BN + 31) / 32;
    if (SWAP) {
      const int feat = n_blk * BLOCK_M + q * 32 + lane;            // N % 128 == 0: always a real feature
      const float bf = (p.bias != nullptr && p.mode != GM_PARTIAL_F32) ? __ldg(p.bias + feat) : 0.f;
      for (int c = half; c < NCHUNK; c += 2) {
        uint32_t r[32];
        if (nkb > 0) {
          tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32, r);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) r[i] = 0u;
        }
        const int m0 = m_blk * BN + c * 32;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int m = m0 + j;
          if (m >= p.M) break;                                      // warp-uniform
          float v = __uint_as_float(r[j]) + bf;
          if (p.mode == GM_PARTIAL_F32) {
            reinterpret_cast<float*>(p.out)[(static_cast<size_t>(z) * p.M + m) * p.ldo + feat] = v;
          } else if (p.mode == GM_STORE_F32) {
            reinterpret_cast<float*>(p.out)[static_cast<size_t>(m) * p.ldo + feat] = v;
          } else {
            if (p.mode == GM_GELU_BF16) v = gelu_erf_fast(v);
            reinterpret_cast<bf16*>(p.out)[static_cast<size_t>(m) * p.ldo + feat] = __float2bfloat16(v);
          }
        }
      }
    } else
    for (int c = half; c < NCHUNK; c += 2) {
      uint32_t r[32];
      if (nkb > 0) {
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32, r);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) r[i] = 0u;
      }
      const int col0 = n_blk * BN + c * 32;
      if (row < p.M) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int col = col0 + g * 8;
          if (col < p.N) epilogue8(p, row, col, z, r + g * 8);
        }
      }
    }
    tc_fence_before();
    if (threadIdx.x == 64) TRACE(5, clock64());
  }
  // No CTA may exit while a peer can still multicast into its shared memory or arrive on its
  // barriers: the whole cluster leaves together.
  if (CL > 1) cluster_sync_all(); else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

// ---------------------------------------------------------------------------
// Persistent variant for the dense launches (many output tiles, short K: the decoder's mid stages and
// transposed convs).  One CTA per SM walks (m, n) tiles round-robin; the TMA ring and its barrier
// phases simply continue from tile to tile and the accumulator is double-buffered in TMEM, so the
// epilogue of tile i (GELU / residual / stores: 3-10 k clk, longer than an 8-k-block main loop) overlaps
// the loads and MMAs of tile i+1.  The one-tile-per-CTA kernel above pays prologue + first-tile TMA
// latency + epilogue serially for every tile (trace r01: 0.7 k + 4.2 k + 3.4 k clk per 128x64 tile).
// ---------------------------------------------------------------------------
template <int BN>
__global__ void __launch_bounds__(NUM_THREADS, 1)
vv_gemm_persistent_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                          const GemmKParams p, int n_tiles_n, int n_items) {
  constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;
  constexpr int B_BYTES = BN * BLOCK_K * 2;
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr int MAXST = 12;
  const int ST = p.stages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + ST * A_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + ST * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + MAXST;
  uint64_t* acc_full = empty_bar + MAXST;      // [2]
  uint64_t* acc_empty = acc_full + 2;          // [2], one arrival per epilogue warp
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < ST; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&acc_full[b], 1);
      mbar_init(&acc_empty[b], NUM_EPI_WARPS);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 2 * BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_launch();
  griddep_wait();                               // activations, residuals and outputs belong to the predecessor until now
  const int kb_total = (p.K + BLOCK_K - 1) / BLOCK_K;
  const int tiles_mn = n_items / max(1, (kb_total + p.kb_per_split - 1) / p.kb_per_split);

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const int z = item / tiles_mn, t = item % tiles_mn;
        const int m_blk = t / n_tiles_n, n_blk = t % n_tiles_n;
        const int kb0 = z * p.kb_per_split, kb1 = min(kb_total, kb0 + p.kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], STAGE_BYTES);
          tma_load_2d(smem_a + stage * A_BYTES, &tmA, &full_bar[stage], kb * BLOCK_K, m_blk * BLOCK_M, p.a_policy);
          tma_load_2d(smem_b + stage * B_BYTES, &tmB, &full_bar[stage], kb * BLOCK_K, n_blk * BN, p.b_policy);
          if (++stage == ST) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    {                                     // converged warp, elect.sync-issued (see vv_gemm_kernel)
      constexpr uint32_t idesc = umma_idesc_bf16(BLOCK_M, BN);
      int stage = 0;
      uint32_t phase = 0, acc_cnt = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++acc_cnt) {
        const int z = item / tiles_mn;
        const int kb0 = z * p.kb_per_split, kb1 = min(kb_total, kb0 + p.kb_per_split);
        const uint32_t buf = acc_cnt & 1;
        mbar_wait(&acc_empty[buf], ((acc_cnt >> 1) & 1) ^ 1);     // the epilogue has drained this buffer
        tc_fence_after();
        const uint32_t d = tmem_base + buf * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint64_t adesc = umma_desc_sw128(smem_u32(smem_a + stage * A_BYTES));
          const uint64_t bdesc = umma_desc_sw128(smem_u32(smem_b + stage * B_BYTES));
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
            umma_bf16_warp(d, adesc + 2 * k, bdesc + 2 * k, idesc, (kb - kb0) | k);
          umma_commit_warp(&empty_bar[stage]);
          if (++stage == ST) { stage = 0; phase ^= 1; }
        }
        umma_commit_warp(&acc_full[buf]);
      }
    }
  } else {
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    uint32_t acc_cnt = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++acc_cnt) {
      const int z = item / tiles_mn, t = item % tiles_mn;
      const int m_blk = t / n_tiles_n, n_blk = t % n_tiles_n;
      const uint32_t buf = acc_cnt & 1;
      mbar_wait(&acc_full[buf], (acc_cnt >> 1) & 1);
      tc_fence_after();
      const int row = m_blk * BLOCK_M + q * 32 + lane;
      constexpr int NCHUNK = BN / 32;
#pragma unroll 1
      for (int c = half; c < NCHUNK; c += 2) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * BN + c * 32, r);
        tmem_ld_wait();
        if (c + 2 >= NCHUNK) {                 // last chunk of this warp: values are in registers, release the buffer
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&acc_empty[buf]);
        }
        const int col0 = n_blk * BN + c * 32;
        if (row < p.M) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const int col = col0 + g * 8;
            if (col < p.N) epilogue8(p, row, col, z, r + g * 8);
          }
        }
      }
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * BN);
  }
}

// ---------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

}  // namespace

// 2-D bf16 tensor [rows, cols] with row stride ld (elements); box = [box_rows, 64], SWIZZLE_128B.
int gemm_make_tmap(CUtensorMap* m, const bf16* ptr, int rows, int cols, int ld, int box_rows) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) {
    vv_set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return VV_ERR_CUDA;
  }
  VV_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "GEMM operand not 16-byte aligned");
  VV_REQUIRE((ld % 8) == 0, "GEMM operand row stride %d not a multiple of 8 elements", ld);
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(BLOCK_K), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<bf16*>(ptr), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    vv_set_error("cuTensorMapEncodeTiled failed (%d) rows=%d cols=%d ld=%d box=%d", (int)r, rows,
                 cols, ld, box_rows);
    return VV_ERR_CUDA;
  }
  return VV_OK;
}

namespace {

template <int BN>
int launch_swapped(const GemmArgs& a, cudaStream_t stream) {
  using Cfg = TileCfg<BN>;
  static bool attr_flags[VV_MAX_DEVICES] = {};
  bool& attr_set = attr_flags[vv_current_device()];   // function attributes are per device
  if (!attr_set) {
    VV_CHECK_CUDA(cudaFuncSetAttribute(vv_gemm_kernel<BN, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_SMEM));
    attr_set = true;
  }
  VV_REQUIRE(a.N % BLOCK_M == 0, "gemm(swapped): N=%d must be a multiple of 128", a.N);
  VV_REQUIRE(a.mode == GM_PARTIAL_F32 || a.mode == GM_STORE_F32 || a.mode == GM_STORE_BF16 || a.mode == GM_GELU_BF16,
             "gemm(swapped): mode %d", a.mode);
  int stages = a.stages;
  while (stages > 1 && Cfg::smem_bytes(stages) > MAX_SMEM) --stages;
  VV_REQUIRE(stages >= 1 && stages <= Cfg::MAX_STAGES, "gemm: stages=%d", stages);
  CUtensorMap tmA, tmB;
  VV_TRY(gemm_make_tmap(&tmA, a.B, a.N, a.K, a.ldb, BLOCK_M));     // weights on the M side
  VV_TRY(gemm_make_tmap(&tmB, a.A, a.M, a.K, a.lda, BN));          // activations on the N side
  GemmKParams p;
  p.M = a.M; p.N = a.N; p.K = a.K; p.mode = a.mode;
  p.out = a.out; p.ldo = a.ldo; p.bias = a.bias; p.gamma = a.gamma;
  p.resid = a.resid; p.ldr = a.ldr; p.gate = a.gate; p.ldg = a.ldg;
  const int kb_total = vv_ceil_div(a.K, BLOCK_K);
  const int splits = a.splits < 1 ? 1 : a.splits;
  p.kb_per_split = vv_ceil_div(kb_total, splits);
  p.stages = stages;
  p.a_policy = VV_L2_EVICT_NORMAL;
  p.b_policy = a.b_streaming ? VV_L2_EVICT_FIRST : VV_L2_EVICT_NORMAL;
  p.a_rows = BLOCK_M;
  dim3 grid(a.N / BLOCK_M, vv_ceil_div(a.M, BN), splits);
  VV_CHECK_CUDA(vv_launch(vv_gemm_kernel<BN, 1, true>, grid, dim3(NUM_THREADS), Cfg::smem_bytes(stages), stream, 1,
                          a.pdl != 0, tmA, tmB, p));
  return VV_OK;
}

template <int BN, int CL>
int launch_bn(const GemmArgs& a, cudaStream_t stream) {
  using Cfg = TileCfg<BN>;
  static bool attr_flags[VV_MAX_DEVICES] = {};
  bool& attr_set = attr_flags[vv_current_device()];   // function attributes are per device
  if (!attr_set) {
    VV_CHECK_CUDA(cudaFuncSetAttribute(vv_gemm_kernel<BN, CL>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_SMEM));
    attr_set = true;
  }
  int stages = a.stages;
  while (stages > 1 && Cfg::smem_bytes(stages) > MAX_SMEM) --stages;
  VV_REQUIRE(stages >= 1 && stages <= Cfg::MAX_STAGES, "gemm: stages=%d", stages);
  CUtensorMap tmA, tmB;
  // skinny problems (decoder stage 0 at n <= 64, single-utterance streaming): transfer only the rows
  // that exist - the SM's TMA ingest, not HBM, bounds these launches and A is most of it
  const int a_rows = (CL == 1 && a.M <= 64) ? (a.M + 7) / 8 * 8 : BLOCK_M;
  VV_TRY(gemm_make_tmap(&tmA, a.A, a.M, a.K, a.lda, CL == 1 ? a_rows : BLOCK_M / CL));
  VV_TRY(gemm_make_tmap(&tmB, a.B, a.N, a.K, a.ldb, BN));
  GemmKParams p;
  p.M = a.M; p.N = a.N; p.K = a.K; p.mode = a.mode;
  p.out = a.out; p.ldo = a.ldo; p.bias = a.bias; p.gamma = a.gamma;
  p.resid = a.resid; p.ldr = a.ldr; p.gate = a.gate; p.ldg = a.ldg;
  const int kb_total = vv_ceil_div(a.K, BLOCK_K);
  const int splits = a.splits < 1 ? 1 : a.splits;
  p.kb_per_split = vv_ceil_div(kb_total, splits);
  p.stages = stages;
  p.a_policy = VV_L2_EVICT_NORMAL;
  p.b_policy = a.b_streaming ? VV_L2_EVICT_FIRST : VV_L2_EVICT_NORMAL;
  p.a_rows = a_rows;
  // the grid is padded to whole clusters; surplus CTAs load zero-filled tiles and store nothing
  dim3 grid(vv_ceil_div(vv_ceil_div(a.N, BN), CL) * CL, vv_ceil_div(a.M, BLOCK_M), splits);
  VV_CHECK_CUDA(vv_launch(vv_gemm_kernel<BN, CL>, grid, dim3(NUM_THREADS), Cfg::smem_bytes(stages), stream, CL,
                          a.pdl != 0, tmA, tmB, p));
  return VV_OK;
}

template <int BN>
int launch_persistent(const GemmArgs& a, cudaStream_t stream) {
  constexpr int STAGE_BYTES = (BLOCK_M + BN) * BLOCK_K * 2;
  constexpr int CTRL_BYTES = 512;
  static bool attr_flags[VV_MAX_DEVICES] = {};
  bool& attr_set = attr_flags[vv_current_device()];   // function attributes are per device
  if (!attr_set) {
    VV_CHECK_CUDA(cudaFuncSetAttribute(vv_gemm_persistent_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_SMEM));
    attr_set = true;
  }
  int stages = (MAX_SMEM - 1024 - CTRL_BYTES) / STAGE_BYTES;
  if (stages > 12) stages = 12;
  CUtensorMap tmA, tmB;
  VV_TRY(gemm_make_tmap(&tmA, a.A, a.M, a.K, a.lda, BLOCK_M));
  VV_TRY(gemm_make_tmap(&tmB, a.B, a.N, a.K, a.ldb, BN));
  GemmKParams p;
  p.M = a.M; p.N = a.N; p.K = a.K; p.mode = a.mode;
  p.out = a.out; p.ldo = a.ldo; p.bias = a.bias; p.gamma = a.gamma;
  p.resid = a.resid; p.ldr = a.ldr; p.gate = a.gate; p.ldg = a.ldg;
  const int kb_total = vv_ceil_div(a.K, BLOCK_K);
  const int splits = a.splits < 1 ? 1 : a.splits;
  p.kb_per_split = vv_ceil_div(kb_total, splits);
  p.stages = stages;
  p.a_policy = VV_L2_EVICT_NORMAL;
  p.b_policy = a.b_streaming ? VV_L2_EVICT_FIRST : VV_L2_EVICT_NORMAL;
  p.a_rows = BLOCK_M;
  const int n_tiles_n = vv_ceil_div(a.N, BN);
  const int n_items = vv_ceil_div(a.M, BLOCK_M) * n_tiles_n * splits;
  const int sms = vv_num_sms();
  const int grid = n_items < sms ? n_items : sms;
  VV_CHECK_CUDA(vv_launch(vv_gemm_persistent_kernel<BN>, dim3(grid), dim3(NUM_THREADS),
                          1024 + stages * STAGE_BYTES + CTRL_BYTES, stream, 1, a.pdl != 0, tmA, tmB, p, n_tiles_n, n_items));
  return VV_OK;
}

template <int BN>
int launch_cl(const GemmArgs& a, cudaStream_t stream) {
  switch (a.cluster) {
    case 1: return launch_bn<BN, 1>(a, stream);
    case 2: return launch_bn<BN, 2>(a, stream);
    case 4: return launch_bn<BN, 4>(a, stream);
    case 8: return launch_bn<BN, 8>(a, stream);
    default:
      vv_set_error("gemm: unsupported cluster size %d", a.cluster);
      return VV_ERR_INVALID;
  }
}

}  // namespace

// Planner (numbers from scripts/sweep_gemm.py on B200, profiles/r01_gemm_sweep.md).
//
// What bounds a launch: an SM ingests at most ~46 B/clk through TMA (measured: a k-block of
// A 16 KB + B 8 KB arrives every 520 clk whatever the ring depth or cluster shape), i.e. ~13 TB/s
// over 148 SMs, twice the HBM rate.  A weight-streaming launch (one M tile) is therefore
// HBM-bound only while the activation bytes an SM pulls stay below its weight bytes and every
// SM is pulling.  Multicasting A across a cluster does not lower the per-SM ingest (each SM
// still receives the whole tile) and measured 3-8 % slower, so the default cluster is 1.
//  * streaming (M <= 128): BN = 64, 4-deep ring (96 KB: two consecutive launches co-reside on
//    an SM, so under PDL the next launch's weight prefetch overlaps this one's epilogue);
//    split-K to ~2 CTAs per SM when the mode allows partial sums.
//  * dense: the widest tile that still leaves >= 4 waves of CTAs, else BN = 64.
void gemm_plan(GemmArgs& a, int num_sms) {
  const int m_tiles = vv_ceil_div(a.M, BLOCK_M);
  const int kb_total = vv_ceil_div(a.K, BLOCK_K);
  const bool streaming = (m_tiles == 1);
  if (a.swap == 0) a.swap = -1;          // callers opt in with swap = 2
  if (a.swap == 2) {
    // swapped operands (see vv_gemm_kernel): split-K partial launches with at most 64 rows - the callers that can sum
    // partials are exactly the ones whose N-tiles alone cannot fill the machine
    a.swap = (vv_option_gemm_swap() && a.mode == GM_PARTIAL_F32 && a.M <= 64 && a.N % BLOCK_M == 0 && a.bn == 0 &&
              a.cluster <= 1 && a.persistent <= 0) ? 1 : -1;
  }
  if (a.swap > 0) {
    if (a.bn == 0) a.bn = (a.M <= 32) ? 32 : 64;
    const int tiles = (a.N / BLOCK_M) * vv_ceil_div(a.M, a.bn);
    if (a.splits == 0) {
      a.splits = 1;
      if (a.mode == GM_PARTIAL_F32) {
        int s = 2 * num_sms / (tiles > 0 ? tiles : 1);
        if (s < 1) s = 1;
        const int max_s = kb_total / 4 > 0 ? kb_total / 4 : 1;
        if (s > max_s) s = max_s;
        if (s > 32) s = 32;
        while (s > 1 && vv_ceil_div(kb_total, s) * (s - 1) >= kb_total) --s;
        a.splits = s;
      }
    }
    a.cluster = 1;
    a.persistent = -1;
    if (a.stages == 0) {
      int st = 4;
      const int kb_per = vv_ceil_div(kb_total, a.splits < 1 ? 1 : a.splits);
      if (st > kb_per) st = kb_per < 1 ? 1 : kb_per;
      a.stages = st;
    }
    return;
  }
  if (a.bn == 0) {
    if (streaming) {
      a.bn = (a.N >= 64) ? 64 : 32;
      // (vv_set_option("gemm_stream_bn32", 0) turns it off) 32-wide tiles when 64-wide ones leave less than two
      // CTAs per SM and K cannot be split - two co-resident CTAs interleave their issue-bound k-loops
      if (vv_option_gemm_stream_bn32() && a.mode != GM_PARTIAL_F32 && a.mode != GM_SWIGLU_BF16 && a.N >= 64 &&
          vv_ceil_div(a.N, 64) < 2 * num_sms && a.N % 32 == 0)
        a.bn = 32;
    } else {
      a.bn = (a.N >= 64) ? 64 : 32;
      const int t128 = m_tiles * vv_ceil_div(a.N, 128), t256 = m_tiles * vv_ceil_div(a.N, 256);
      if (a.persistent >= 0 && a.cluster <= 1 && a.N % 128 == 0 && t128 > num_sms && vv_option_gemm_persistent()) {
        // more than a wave of 128-wide tiles: the persistent kernel hides every epilogue but the last, so
        // wide tiles win (one_gemm sweep, r01: 2560x2048x512 18.9 -> 15.9 us, 12800x512x512 18.6 -> 13.6,
        // 51200x256x256 26.3 -> 16.6).  256 columns only when that does not cost an extra half wave.
        const int w128 = vv_ceil_div(t128, num_sms), w256 = vv_ceil_div(t256, num_sms);
        a.bn = (a.N % 256 == 0 && 2 * w256 <= w128) ? 256 : 128;
        a.persistent = 1;
      } else {
        const int cands[2] = {256, 128};
        const int tl[2] = {t256, t128};
        for (int i = 0; i < 2; ++i)
          if (cands[i] <= a.N && tl[i] >= 4 * num_sms) { a.bn = cands[i]; break; }
        if (a.persistent == 0) a.persistent = -1;
      }
    }
  }
  const int n_tiles = vv_ceil_div(a.N, a.bn);
  if (a.splits == 0) {
    a.splits = 1;
    if (a.mode == GM_PARTIAL_F32) {
      const int tiles = m_tiles * n_tiles;
      int s = (streaming ? 2 * num_sms : num_sms) / (tiles > 0 ? tiles : 1);
      if (s < 1) s = 1;
      const int max_s = kb_total / 4 > 0 ? kb_total / 4 : 1;   // >= 4 k-blocks per split
      if (s > max_s) s = max_s;
      if (s > 32) s = 32;                                      // scratch is sized for <= 32 splits
      // make sure every split owns at least one k-block
      while (s > 1 && vv_ceil_div(kb_total, s) * (s - 1) >= kb_total) --s;
      a.splits = s;
    }
  }
  if (a.cluster == 0) a.cluster = 1;
  if (a.persistent == 0) {
    // dense launches with at least a wave of tiles: walk them with one CTA per SM so that every epilogue
    // overlaps the next tile's main loop
    // (only reached with a caller-chosen tile width: the automatic choice above already decided)
    const bool dense = !streaming && m_tiles * n_tiles * (a.splits < 1 ? 1 : a.splits) >= num_sms;
    a.persistent = (dense && a.cluster == 1 && a.bn >= 64 && vv_option_gemm_persistent()) ? 1 : -1;
  }
  if (a.stages == 0) {
    int st = 4;
    const int kb_per = vv_ceil_div(kb_total, a.splits < 1 ? 1 : a.splits);
    if (st > kb_per) st = kb_per < 1 ? 1 : kb_per;
    a.stages = st;
  }
}

int gemm_debug_trace(long long* out, int cap) {
#ifdef VV_GEMM_TRACE
  const int n = 160 * 48;
  if (cap < n) return -1;
  if (cudaMemcpyFromSymbol(out, g_gemm_trace, sizeof(long long) * n) != cudaSuccess) return -2;
  return n;
#else
  (void)out; (void)cap;
  return 0;
#endif
}

size_t gemm_partial_elems(const GemmArgs& a) {
  return static_cast<size_t>(a.splits < 1 ? 1 : a.splits) * a.M * a.ldo;
}

namespace {
struct TimedLaunch { GemmLaunchRec rec; cudaEvent_t e0, e1; };
bool g_timing = false;
std::vector<TimedLaunch> g_timed;
int gemm_launch_impl(const GemmArgs& a, cudaStream_t stream);
}  // namespace

void gemm_timing_enable(bool on) { g_timing = on; }
bool gemm_timing_on() { return g_timing; }
void gemm_timing_push(const GemmLaunchRec& rec, cudaEvent_t e0, cudaEvent_t e1) { g_timed.push_back(TimedLaunch{rec, e0, e1}); }

int gemm_timing_fetch(GemmLaunchRec* out, int cap) {
  int n = 0;
  for (TimedLaunch& t : g_timed) {
    cudaEventSynchronize(t.e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, t.e0, t.e1);
    t.rec.ms = ms;
    if (out && n < cap) out[n++] = t.rec;
    cudaEventDestroy(t.e0);
    cudaEventDestroy(t.e1);
  }
  g_timed.clear();
  return n;
}

int gemm_launch(const GemmArgs& a, cudaStream_t stream) {
  if (!g_timing) return gemm_launch_impl(a, stream);
  TimedLaunch t;
  t.rec = GemmLaunchRec{a.M, a.N, a.K, a.bn, a.splits, a.mode, 0.f, 0, 0};
  VV_CHECK_CUDA(cudaEventCreate(&t.e0));
  VV_CHECK_CUDA(cudaEventCreate(&t.e1));
  VV_CHECK_CUDA(cudaEventRecord(t.e0, stream));
  const int rc = gemm_launch_impl(a, stream);
  VV_CHECK_CUDA(cudaEventRecord(t.e1, stream));
  g_timed.push_back(t);
  return rc;
}

namespace {
int gemm_launch_impl(const GemmArgs& a, cudaStream_t stream) {
  VV_REQUIRE(a.A && a.B && a.out, "gemm: null operand");
  VV_REQUIRE(a.M > 0 && a.N > 0 && a.K > 0, "gemm: empty problem %dx%dx%d", a.M, a.N, a.K);
  VV_REQUIRE(a.N % 8 == 0 && a.K % 8 == 0, "gemm: N=%d and K=%d must be multiples of 8", a.N, a.K);
  VV_REQUIRE(a.mode != GM_SWIGLU_BF16 || (a.ldo % 4 == 0), "gemm: swiglu ldo %% 4");
  VV_REQUIRE(a.mode == GM_SWIGLU_BF16 || a.ldo % 8 == 0, "gemm: ldo=%d must be a multiple of 8", a.ldo);
  VV_REQUIRE(a.splits <= 1 || a.mode == GM_PARTIAL_F32, "gemm: only GM_PARTIAL_F32 may split K");
  VV_REQUIRE(a.mode != GM_RESID_GAMMA_BF16 || (a.resid && a.gamma && a.ldr % 8 == 0),
             "gemm: resid/gamma missing");
  VV_REQUIRE(a.mode != GM_RESID_GATE_F32 || (a.resid && a.gate && a.ldr % 4 == 0 && a.ldg % 8 == 0),
             "gemm: resid/gate missing");
  if (a.swap > 0) {
    switch (a.bn) {
      case 64: return launch_swapped<64>(a, stream);
      case 32: return launch_swapped<32>(a, stream);
      default:
        vv_set_error("gemm(swapped): unsupported row-tile width %d", a.bn);
        return VV_ERR_INVALID;
    }
  }
  if (a.persistent > 0) {
    switch (a.bn) {
      case 256: return launch_persistent<256>(a, stream);
      case 128: return launch_persistent<128>(a, stream);
      case 64: return launch_persistent<64>(a, stream);
      default: break;       // bn = 32: the one-tile kernel below
    }
  }
  switch (a.bn) {
    case 256: return launch_cl<256>(a, stream);
    case 128: return launch_cl<128>(a, stream);
    case 64: return launch_cl<64>(a, stream);
    case 32: return launch_cl<32>(a, stream);
    default:
      vv_set_error("gemm: unsupported tile width %d", a.bn);
      return VV_ERR_INVALID;
  }
}
}  // namespace
