// vv_noalloc.cu — vv_linear_splitk_static: a split-K linear that allocates nothing.
//
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
//
//   out[M,N] = A[M,K] . W[N,K]^T + bias          A, W bf16 row-major; out bf16 or fp32
//
// K is split over CTAs (the weight-streaming shape of this library: few rows against a long K, where N-tiles alone
// cannot fill 148 SMs); each split writes an fp32 partial tile and a second kernel sums them row-wise.  Everywhere
// else in the library that partial-sum scratch is a cudaMalloc'ed buffer owned by a handle.  Here it is a file-scope
// device array, so the entry point calls no cudaMalloc / cudaFree / allocating routine at all:
//   * 4 MiB (VV_SPLITK_STATIC_MAX_FLOATS fp32) of every GPU that loads libvvb200.so, held until the process exits;
//   * splits * M * N must fit - larger problems are rejected with VV_ERR_INVALID and a message;
//   * ONE array per GPU shared by every call: it keeps the previous call's partials, so the region a call uses is
//     cleared first (cudaMemsetAsync on the caller's stream), and two calls that overlap on the GPU (different
//     streams, no dependency between them) race on it - serialise them;
//   * in host code the array's NAME is a host-side stand-in, not a device address: the address comes from
//     cudaGetSymbolAddress, looked up once per device with that device current, below.
#include "../../include/vvb200.h"
#include "vv_common.cuh"
#include "vv_gemm.cuh"

static_assert(VV_SPLITK_STATIC_MAX_FLOATS == (1 << 20), "scratch size is part of the ABI (include/vvb200.h)");

static __device__ __align__(256) float g_splitk_static_scratch[VV_SPLITK_STATIC_MAX_FLOATS];

namespace {

constexpr int NA_MAX_DEVICES = 64;
float* g_scratch_addr[NA_MAX_DEVICES] = {};   // device address of the array on each device (set once)

// Device address of the scratch array on the CURRENT device.
int static_scratch(float** out) {
  int dev = -1;
  VV_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= NA_MAX_DEVICES) {
    vv_set_error("vv_linear_splitk_static: device ordinal %d not supported", dev);
    return VV_ERR_UNSUPPORTED;
  }
  if (!g_scratch_addr[dev]) {
    float* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_splitk_static_scratch);
    if (e != cudaSuccess || !p) {
      vv_set_error("vv_linear_splitk_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return VV_ERR_CUDA;
    }
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return VV_OK;
}

// out[m, n] = sum_z part[z][m][n] + bias[n]      (8 columns per thread; fixed summation order z = 0..nsplit-1)
template <typename T>
__global__ void k_static_reduce(const float* __restrict__ part, int nsplit, const float* __restrict__ bias,
                                T* __restrict__ out, int M, int N, int ldo) {
  griddep_wait();
  griddep_launch();
  const int vpr = N / 8;
  const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  if (i >= static_cast<size_t>(M) * vpr) return;
  const int c0 = static_cast<int>(i % vpr) * 8;
  const size_t m = i / vpr;
  float s[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s[k] = bias ? bias[c0 + k] : 0.f;
  for (int z = 0; z < nsplit; ++z) {
    const float4* p = reinterpret_cast<const float4*>(part + (static_cast<size_t>(z) * M + m) * N + c0);
    const float4 a = p[0], c = p[1];
    s[0] += a.x; s[1] += a.y; s[2] += a.z; s[3] += a.w; s[4] += c.x; s[5] += c.y; s[6] += c.z; s[7] += c.w;
  }
  T* o = out + m * static_cast<size_t>(ldo) + c0;
  if constexpr (sizeof(T) == 2) {
    *reinterpret_cast<uint4*>(o) = pack_bf16x8(s);
  } else {
    reinterpret_cast<float4*>(o)[0] = make_float4(s[0], s[1], s[2], s[3]);
    reinterpret_cast<float4*>(o)[1] = make_float4(s[4], s[5], s[6], s[7]);
  }
}

}  // namespace

extern "C" {

int vv_linear_splitk_static(const void* A, int lda, const void* W, int ldw, int M, int N, int K, int splits,
                            const float* bias, void* out, int ldo, int out_dtype, void* stream) {
  VV_REQUIRE(A && W && out, "vv_linear_splitk_static: null operand");
  VV_REQUIRE(M > 0 && N > 0 && K > 0, "vv_linear_splitk_static: empty problem %dx%dx%d", M, N, K);
  VV_REQUIRE(N % 8 == 0 && K % 8 == 0, "vv_linear_splitk_static: N=%d and K=%d must be multiples of 8", N, K);
  VV_REQUIRE(out_dtype == VV_DTYPE_BF16 || out_dtype == VV_DTYPE_F32, "vv_linear_splitk_static: out_dtype %d", out_dtype);
  VV_REQUIRE(ldo >= N && ldo % (out_dtype == VV_DTYPE_BF16 ? 8 : 4) == 0,
             "vv_linear_splitk_static: ldo=%d (need >= N and 16-byte aligned rows)", ldo);
  VV_REQUIRE(splits >= 0 && splits <= 32, "vv_linear_splitk_static: splits=%d (0 = choose, at most 32)", splits);
  // per-launch timing brackets every GEMM with freshly created events: an allocating routine, not allowed here
  VV_REQUIRE(!gemm_timing_on(), "vv_linear_splitk_static: vv_timing_enable(1) is active; disable it (it creates CUDA events)");
  cudaStream_t st = static_cast<cudaStream_t>(stream);

  GemmArgs g;
  g.A = static_cast<const bf16*>(A); g.lda = lda;
  g.B = static_cast<const bf16*>(W); g.ldb = ldw;
  g.M = M; g.N = N; g.K = K;
  g.mode = GM_PARTIAL_F32; g.ldo = N; g.splits = splits;
  g.b_streaming = (M <= 256);
  gemm_plan(g, vv_num_sms());
  int sp = g.splits < 1 ? 1 : g.splits;
  const int kb = vv_ceil_div(K, 64);
  if (sp > kb) sp = kb;
  while (sp > 1 && vv_ceil_div(kb, sp) * (sp - 1) >= kb) --sp;   // no empty split
  const long long per_split = static_cast<long long>(M) * N;
  if (splits == 0)                                              // chosen here: shrink until the scratch holds it
    while (sp > 1 && per_split * sp > VV_SPLITK_STATIC_MAX_FLOATS) {
      --sp;
      while (sp > 1 && vv_ceil_div(kb, sp) * (sp - 1) >= kb) --sp;
    }
  if (per_split * sp > VV_SPLITK_STATIC_MAX_FLOATS) {
    vv_set_error("vv_linear_splitk_static: splits*M*N = %d*%d*%d = %lld floats exceeds the static scratch of %d floats "
                 "(%d MiB per GPU); use vv_linear with a caller-owned buffer for larger problems",
                 sp, M, N, per_split * sp, VV_SPLITK_STATIC_MAX_FLOATS, VV_SPLITK_STATIC_MAX_FLOATS / (1 << 18));
    return VV_ERR_INVALID;
  }
  g.splits = sp;

  float* scratch = nullptr;
  VV_TRY(static_scratch(&scratch));
  // the array keeps the previous call's partials: clear the region this call sums over
  VV_CHECK_CUDA(cudaMemsetAsync(scratch, 0, static_cast<size_t>(per_split) * sp * sizeof(float), st));
  g.out = scratch;
  g.pdl = 0;                                                    // the memset precedes it: plain stream order
  VV_TRY(gemm_launch(g, st));
  const unsigned grid = static_cast<unsigned>((static_cast<size_t>(M) * (N / 8) + 255) / 256);
  if (out_dtype == VV_DTYPE_BF16)
    VV_CHECK_CUDA(vv_launch(k_static_reduce<bf16>, dim3(grid), dim3(256), 0, st, 1, false,
                            static_cast<const float*>(scratch), sp, bias, static_cast<bf16*>(out), M, N, ldo));
  else
    VV_CHECK_CUDA(vv_launch(k_static_reduce<float>, dim3(grid), dim3(256), 0, st, 1, false,
                            static_cast<const float*>(scratch), sp, bias, static_cast<float*>(out), M, N, ldo));
  return VV_OK;
}

}  // extern "C"
