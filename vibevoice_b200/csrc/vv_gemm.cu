// vv_gemm.cu — tcgen05 / TMEM / TMA GEMM with fused epilogues (see vv_gemm.cuh).
#include "vv_gemm.cuh"

#include <mutex>
#include <vector>

namespace {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;   // 64 bf16 = 128 B = one SWIZZLE_128B row
constexpr int UMMA_K = 16;
constexpr int NUM_EPI_WARPS = 8;
constexpr int NUM_THREADS = 32 * (2 + NUM_EPI_WARPS);

struct GemmKParams {
  int M, N, K;
  int mode;
  void* out; int ldo;
  const float* bias;
  const float* gamma;
  const void* resid; int ldr;
  const bf16* gate; int ldg;
  int kb_per_split;   // k-blocks handled by one blockIdx.z
  unsigned long long a_policy, b_policy;
};

template <int BN>
struct TileCfg {
  static constexpr int STAGES = (BN == 256) ? 4 : (BN == 128) ? 3 : (BN == 64) ? 8 : 5;
  static constexpr int CTAS_PER_SM = (BN == 128 || BN == 32) ? 2 : 1;
  static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;
  static constexpr int B_BYTES = BN * BLOCK_K * 2;
  static constexpr int TMEM_COLS = BN < 32 ? 32 : BN;
  static constexpr int SMEM_BYTES = 1024 + STAGES * (A_BYTES + B_BYTES) + 256;
};

// Epilogue for 8 consecutive accumulator columns of one row.
__device__ __forceinline__ void epilogue8(const GemmKParams& p, int row, int col, int z,
                                          const uint32_t* r) {
  float v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
  const int mode = p.mode;
  if (mode == GM_PARTIAL_F32) {
    float* o = reinterpret_cast<float*>(p.out) + (static_cast<size_t>(z) * p.M + row) * p.ldo + col;
    reinterpret_cast<float4*>(o)[0] = make_float4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<float4*>(o)[1] = make_float4(v[4], v[5], v[6], v[7]);
    return;
  }
  if (mode == GM_RESID_GATE_F32) {
    const float* x = reinterpret_cast<const float*>(p.resid) + static_cast<size_t>(row) * p.ldr + col;
    float4 x0 = reinterpret_cast<const float4*>(x)[0], x1 = reinterpret_cast<const float4*>(x)[1];
    uint4 gv = *reinterpret_cast<const uint4*>(p.gate + static_cast<size_t>(row) * p.ldg + col);
    float g[8];
    unpack_bf16x8(gv, g);
    float* o = reinterpret_cast<float*>(p.out) + static_cast<size_t>(row) * p.ldo + col;
    reinterpret_cast<float4*>(o)[0] = make_float4(x0.x + g[0] * v[0], x0.y + g[1] * v[1],
                                                  x0.z + g[2] * v[2], x0.w + g[3] * v[3]);
    reinterpret_cast<float4*>(o)[1] = make_float4(x1.x + g[4] * v[4], x1.y + g[5] * v[5],
                                                  x1.z + g[6] * v[6], x1.w + g[7] * v[7]);
    return;
  }
  if (p.bias != nullptr) {
    float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + col));
    float4 b1 = __ldg(reinterpret_cast<const float4*>(p.bias + col) + 1);
    v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
    v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
  }
  if (mode == GM_STORE_F32) {
    float* o = reinterpret_cast<float*>(p.out) + static_cast<size_t>(row) * p.ldo + col;
    reinterpret_cast<float4*>(o)[0] = make_float4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<float4*>(o)[1] = make_float4(v[4], v[5], v[6], v[7]);
    return;
  }
  if (mode == GM_SWIGLU_BF16) {
    bf16* o = reinterpret_cast<bf16*>(p.out) + static_cast<size_t>(row) * p.ldo + (col >> 1);
    uint2 w;
    w.x = pack_bf16x2(silu_f(v[0]) * v[1], silu_f(v[2]) * v[3]);
    w.y = pack_bf16x2(silu_f(v[4]) * v[5], silu_f(v[6]) * v[7]);
    *reinterpret_cast<uint2*>(o) = w;
    return;
  }
  if (mode == GM_GELU_BF16) {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = gelu_erf_f(v[i]);
  } else if (mode == GM_SILU_BF16) {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = silu_f(v[i]);
  } else if (mode == GM_RESID_GAMMA_BF16) {
    float4 g0 = __ldg(reinterpret_cast<const float4*>(p.gamma + col));
    float4 g1 = __ldg(reinterpret_cast<const float4*>(p.gamma + col) + 1);
    uint4 xv = *reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(p.resid) +
                                               static_cast<size_t>(row) * p.ldr + col);
    float x[8];
    unpack_bf16x8(xv, x);
    v[0] = x[0] + g0.x * v[0]; v[1] = x[1] + g0.y * v[1];
    v[2] = x[2] + g0.z * v[2]; v[3] = x[3] + g0.w * v[3];
    v[4] = x[4] + g1.x * v[4]; v[5] = x[5] + g1.y * v[5];
    v[6] = x[6] + g1.z * v[6]; v[7] = x[7] + g1.w * v[7];
  }
  bf16* o = reinterpret_cast<bf16*>(p.out) + static_cast<size_t>(row) * p.ldo + col;
  *reinterpret_cast<uint4*>(o) = pack_bf16x8(v);
}

template <int BN>
__global__ void __launch_bounds__(NUM_THREADS, TileCfg<BN>::CTAS_PER_SM)
vv_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const GemmKParams p) {
  using Cfg = TileCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B operands need 1024-byte aligned tiles.
  uint8_t* smem = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * Cfg::A_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * (Cfg::A_BYTES + Cfg::B_BYTES));
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* accum_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n_blk = blockIdx.x, m_blk = blockIdx.y, z = blockIdx.z;

  const int kb_total = (p.K + BLOCK_K - 1) / BLOCK_K;
  const int kb0 = z * p.kb_per_split;
  const int kb1 = min(kb_total, kb0 + p.kb_per_split);
  const int nkb = max(kb1 - kb0, 0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(accum_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        mbar_arrive_expect_tx(&full_bar[stage], Cfg::A_BYTES + Cfg::B_BYTES);
        tma_load_2d(smem_a + stage * Cfg::A_BYTES, &tmA, &full_bar[stage], kb * BLOCK_K,
                    m_blk * BLOCK_M, p.a_policy);
        tma_load_2d(smem_b + stage * Cfg::B_BYTES, &tmB, &full_bar[stage], kb * BLOCK_K,
                    n_blk * BN, p.b_policy);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer --------------------------------
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(BLOCK_M, BN);
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint64_t adesc = umma_desc_sw128(smem_u32(smem_a + stage * Cfg::A_BYTES));
        const uint64_t bdesc = umma_desc_sw128(smem_u32(smem_b + stage * Cfg::B_BYTES));
#pragma unroll
        for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
          // advance 32 B (= UMMA_K bf16) inside the 128-B swizzle row: +2 in the >>4 field
          umma_bf16(tmem_base, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
        }
        umma_commit(&empty_bar[stage]);   // frees the smem slot when these MMAs retire
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      umma_commit(accum_bar);             // accumulator complete
    }
  } else {
    // ------------------------------ epilogue ----------------------------------
    const int q = warp & 3;               // TMEM lane quadrant this warp may read
    const int half = (warp - 2) >> 2;     // two warps share a quadrant; split the columns
    mbar_wait(accum_bar, 0);
    tc_fence_after();
    const int row = m_blk * BLOCK_M + q * 32 + lane;
    constexpr int NCHUNK = (BN + 31) / 32;
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
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
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

// 2-D bf16 tensor [rows, cols] with row stride ld (elements); box = [box_rows, 64].
int make_tmap(CUtensorMap* m, const bf16* ptr, int rows, int cols, int ld, int box_rows) {
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

template <int BN>
int launch_bn(const GemmArgs& a, cudaStream_t stream) {
  using Cfg = TileCfg<BN>;
  static bool attr_set = false;
  if (!attr_set) {
    VV_CHECK_CUDA(cudaFuncSetAttribute(vv_gemm_kernel<BN>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg::SMEM_BYTES));
    attr_set = true;
  }
  CUtensorMap tmA, tmB;
  VV_TRY(make_tmap(&tmA, a.A, a.M, a.K, a.lda, BLOCK_M));
  VV_TRY(make_tmap(&tmB, a.B, a.N, a.K, a.ldb, BN));
  GemmKParams p;
  p.M = a.M; p.N = a.N; p.K = a.K; p.mode = a.mode;
  p.out = a.out; p.ldo = a.ldo; p.bias = a.bias; p.gamma = a.gamma;
  p.resid = a.resid; p.ldr = a.ldr; p.gate = a.gate; p.ldg = a.ldg;
  const int kb_total = vv_ceil_div(a.K, BLOCK_K);
  const int splits = a.splits < 1 ? 1 : a.splits;
  p.kb_per_split = vv_ceil_div(kb_total, splits);
  p.a_policy = VV_L2_EVICT_NORMAL;
  p.b_policy = a.b_streaming ? VV_L2_EVICT_FIRST : VV_L2_EVICT_NORMAL;
  dim3 grid(vv_ceil_div(a.N, BN), vv_ceil_div(a.M, BLOCK_M), splits);
  vv_gemm_kernel<BN><<<grid, NUM_THREADS, Cfg::SMEM_BYTES, stream>>>(tmA, tmB, p);
  VV_COUNT_LAUNCH();
  VV_CHECK_CUDA(cudaGetLastError());
  return VV_OK;
}

}  // namespace

void gemm_plan(GemmArgs& a, int num_sms) {
  const int m_tiles = vv_ceil_div(a.M, BLOCK_M);
  if (a.bn == 0) {
    const int cands[4] = {256, 128, 64, 32};
    int pick = 32;
    for (int i = 0; i < 4; ++i) {
      const int bn = cands[i];
      if (bn > 32 && bn > a.N) continue;          // don't use a tile wider than the problem
      if (m_tiles * vv_ceil_div(a.N, bn) >= num_sms) { pick = bn; break; }
      pick = bn;                                   // fall through to the narrowest usable tile
    }
    // Small grids: prefer the tile that yields the most CTAs but keep >= 64 columns when the
    // problem is wide, so the A operand is not re-fetched by too many CTAs.
    if (m_tiles * vv_ceil_div(a.N, pick) < num_sms && pick < 64 && a.N >= 64 * num_sms / 2)
      pick = 64;
    a.bn = pick;
  }
  if (a.splits == 0) {
    a.splits = 1;
    if (a.mode == GM_PARTIAL_F32) {
      const int tiles = m_tiles * vv_ceil_div(a.N, a.bn);
      const int kb_total = vv_ceil_div(a.K, BLOCK_K);
      int s = num_sms / (tiles > 0 ? tiles : 1);
      if (s < 1) s = 1;
      const int max_s = kb_total / 4 > 0 ? kb_total / 4 : 1;   // >= 4 k-blocks per split
      if (s > max_s) s = max_s;
      if (s > 32) s = 32;                                      // scratch is sized for <= 32 splits
      // make sure every split owns at least one k-block
      while (s > 1 && vv_ceil_div(kb_total, s) * (s - 1) >= kb_total) --s;
      a.splits = s;
    }
  }
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
  t.rec = GemmLaunchRec{a.M, a.N, a.K, a.bn, a.splits, a.mode, 0.f};
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
  switch (a.bn) {
    case 256: return launch_bn<256>(a, stream);
    case 128: return launch_bn<128>(a, stream);
    case 64: return launch_bn<64>(a, stream);
    case 32: return launch_bn<32>(a, stream);
    default:
      vv_set_error("gemm: unsupported tile width %d", a.bn);
      return VV_ERR_INVALID;
  }
}
}  // namespace
