// vv_gemm_dev.cuh — device-side pieces shared by the GEMM kernels (vv_gemm.cu) and the persistent
// phase engine (vv_chain.cu): tile constants, kernel parameter block, fused epilogues.
#pragma once
#include "vv_gemm.cuh"

namespace {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;   // 64 bf16 = 128 B = one SWIZZLE_128B row
constexpr int UMMA_K = 16;
constexpr int NUM_EPI_WARPS = 8;
constexpr int NUM_THREADS = 32 * (2 + NUM_EPI_WARPS);
constexpr int MAX_SMEM = 227 * 1024;   // opt-in dynamic shared memory per CTA on sm_100

struct GemmKParams {
  int M, N, K;
  int mode;
  void* out; int ldo;
  const float* bias;
  const float* gamma;
  const void* resid; int ldr;
  const bf16* gate; int ldg;
  int kb_per_split;   // k-blocks handled by one blockIdx.z
  int stages;         // depth of the shared-memory ring (runtime: sized by the planner)
  unsigned long long a_policy, b_policy;
  int a_rows;         // rows of the A box TMA transfers per k-block (128, or M rounded up to 8 when M <= 64:
                      // the rest of the 128-row smem tile is never read back - rows >= M are not stored)
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
    for (int i = 0; i < 8; i += 2) {
      const float2 o = gelu_erf_fast2(make_float2(v[i], v[i + 1]));
      v[i] = o.x;
      v[i + 1] = o.y;
    }
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

}  // namespace
