// vv_conn.cu — SpeechConnector: fc1 -> LlamaRMSNorm(eps 1e-6) -> fc2
// (modeling_vibevoice.py:58-69), applied per frame to the acoustic latent [n,1,64] and to the semantic
// features [n,1,128]; the two results are summed into the next LLM input embedding
// (modeling_vibevoice_inference.py:682-687).
//
// Three launches per connector, all on the caller's stream: fc1 as a tcgen05 GEMM with an fp32 + bias
// epilogue, one row-wise RMSNorm kernel, fc2 as a GEMM whose epilogue adds the bias and, when `addend` is
// given, the other connector's output (so acoustic + semantic needs no extra pass).
#include <algorithm>

#include "../../include/vvb200.h"
#include "vv_common.cuh"
#include "vv_gemm.cuh"

namespace {

// y[m, :] = bf16( bf16(x[m, :] * rsqrt(mean(x^2) + eps)) * w )   (LlamaRMSNorm: fp32 statistics, cast, then weight)
__global__ void __launch_bounds__(256) k_conn_norm(const float* __restrict__ x, const float* __restrict__ w,
                                                   bf16* __restrict__ y, int H, float eps) {
  griddep_wait();
  griddep_launch();
  __shared__ float red[8];
  const int m = blockIdx.x, tid = threadIdx.x;
  const float4* xr = reinterpret_cast<const float4*>(x + static_cast<size_t>(m) * H);
  float ss = 0.f;
  for (int i = tid; i < H / 4; i += 256) {
    const float4 v = xr[i];
    ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  ss = warp_sum(ss);
  if ((tid & 31) == 0) red[tid >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) tot += red[i];
  const float rms = rsqrtf(tot / static_cast<float>(H) + eps);
  for (int i = tid; i < H / 4; i += 256) {
    const float4 v = xr[i];
    const float4 g = __ldg(reinterpret_cast<const float4*>(w) + i);
    const float a = __bfloat162float(__float2bfloat16(v.x * rms)) * g.x, b = __bfloat162float(__float2bfloat16(v.y * rms)) * g.y;
    const float c = __bfloat162float(__float2bfloat16(v.z * rms)) * g.z, d = __bfloat162float(__float2bfloat16(v.w * rms)) * g.w;
    uint2 o;
    o.x = pack_bf16x2(a, b);
    o.y = pack_bf16x2(c, d);
    *reinterpret_cast<uint2*>(y + static_cast<size_t>(m) * H + i * 4) = o;
  }
}

template <typename T>
__global__ void k_to_bf16(const T* __restrict__ in, bf16* __restrict__ out, size_t n) {
  griddep_wait();
  griddep_launch();
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n; i += static_cast<size_t>(gridDim.x) * blockDim.x)
    out[i] = __float2bfloat16(static_cast<float>(in[i]));
}
template <typename T>
__global__ void k_to_f32(const T* __restrict__ in, float* __restrict__ out, size_t n) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n; i += static_cast<size_t>(gridDim.x) * blockDim.x)
    out[i] = static_cast<float>(in[i]);
}
__global__ void k_fill_f32(float* out, size_t n, float v) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n; i += static_cast<size_t>(gridDim.x) * blockDim.x) out[i] = v;
}

inline dim3 g1(size_t n, int block) { return dim3(static_cast<unsigned>((n + block - 1) / block)); }

// Voice-prompt latents: VibeVoiceTokenizerEncoderOutput.sample (:1018-1046) fused with the
// speech bias / scaling of _process_speech_inputs (modeling_vibevoice_inference.py:156-159).  Every op rounds to the
// io dtype like the eager reference does (no contraction into FMAs):
//   s   = row_noise ? T(row_noise[b] * std) : std           ('gaussian': one std per utterance; 'fix': the scalar)
//   out = T(T(T(mean + T(s * noise)) + bias) * scale)        (noise == nullptr: out = T(T(mean + bias) * scale))
template <typename T> __device__ __forceinline__ float rnd(float v);
template <> __device__ __forceinline__ float rnd<float>(float v) { return v; }
template <> __device__ __forceinline__ float rnd<bf16>(float v) { return __bfloat162float(__float2bfloat16(v)); }
template <typename T> __device__ __forceinline__ T cvt_out(float v);
template <> __device__ __forceinline__ float cvt_out<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 cvt_out<bf16>(float v) { return __float2bfloat16(v); }

template <typename T>
__global__ void k_latent_sample(const T* __restrict__ mean, const T* __restrict__ noise, const T* __restrict__ row_noise,
                                float std, float bias, float scale, size_t per_row, size_t total, T* __restrict__ out,
                                T* __restrict__ std_out) {
  griddep_wait();
  griddep_launch();
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t b = i / per_row;
    float v = static_cast<float>(mean[i]);
    if (noise != nullptr) {
      float sd = std;
      if (row_noise != nullptr) {
        sd = rnd<T>(__fmul_rn(static_cast<float>(row_noise[b]), std));
        if (std_out != nullptr && i == b * per_row) std_out[b] = cvt_out<T>(sd);
      }
      v = rnd<T>(__fadd_rn(v, rnd<T>(__fmul_rn(sd, static_cast<float>(noise[i])))));
    }
    v = rnd<T>(__fadd_rn(v, bias));
    out[i] = cvt_out<T>(__fmul_rn(v, scale));
  }
}

}  // namespace

struct VVConn {
  int in_dim = 0, out_dim = 0, max_rows = 0;
  float eps = 1e-6f;
  DevBuf w1, b1, nw, w2, b2, ones, xin, h, y;
};

extern "C" {

int vv_conn_create(int in_dim, int out_dim, float eps, int max_rows, int weight_dtype, const void* const* weights,
                   int n_weights, VVConn** out) {
  VV_REQUIRE(weights && out, "vv_conn_create: null argument");
  VV_REQUIRE(n_weights == 5, "vv_conn_create: expected 5 weights (fc1.weight, fc1.bias, norm.weight, fc2.weight, fc2.bias), got %d", n_weights);
  VV_REQUIRE(in_dim >= 8 && in_dim % 8 == 0 && out_dim >= 8 && out_dim % 8 == 0, "vv_conn_create: dims %d -> %d must be multiples of 8", in_dim, out_dim);
  VV_REQUIRE(max_rows >= 1, "vv_conn_create: max_rows=%d", max_rows);
  VV_REQUIRE(weight_dtype == VV_DTYPE_BF16 || weight_dtype == VV_DTYPE_F32, "vv_conn_create: weight_dtype");
  for (int i = 0; i < 5; ++i) VV_REQUIRE(weights[i] != nullptr, "vv_conn_create: weight %d is null", i);
  VVConn* c = new VVConn();
  c->in_dim = in_dim; c->out_dim = out_dim; c->eps = eps; c->max_rows = max_rows;
  const size_t H = out_dim, I = in_dim, R = max_rows;
  int rc = c->w1.alloc(H * I * 2);
  if (rc == VV_OK) rc = c->b1.alloc(H * 4);
  if (rc == VV_OK) rc = c->nw.alloc(H * 4);
  if (rc == VV_OK) rc = c->w2.alloc(H * H * 2);
  if (rc == VV_OK) rc = c->b2.alloc(H * 4);
  if (rc == VV_OK) rc = c->ones.alloc(H * 4);
  if (rc == VV_OK) rc = c->xin.alloc(R * I * 2);
  if (rc == VV_OK) rc = c->h.alloc(R * H * 4);
  if (rc == VV_OK) rc = c->y.alloc(R * H * 2);
  if (rc != VV_OK) { delete c; return rc; }
  if (weight_dtype == VV_DTYPE_F32) {
    k_to_bf16<float><<<256, 256>>>(static_cast<const float*>(weights[0]), c->w1.as<bf16>(), H * I);
    k_to_f32<float><<<64, 256>>>(static_cast<const float*>(weights[1]), c->b1.as<float>(), H);
    k_to_f32<float><<<64, 256>>>(static_cast<const float*>(weights[2]), c->nw.as<float>(), H);
    k_to_bf16<float><<<1024, 256>>>(static_cast<const float*>(weights[3]), c->w2.as<bf16>(), H * H);
    k_to_f32<float><<<64, 256>>>(static_cast<const float*>(weights[4]), c->b2.as<float>(), H);
  } else {
    k_to_bf16<bf16><<<256, 256>>>(static_cast<const bf16*>(weights[0]), c->w1.as<bf16>(), H * I);
    k_to_f32<bf16><<<64, 256>>>(static_cast<const bf16*>(weights[1]), c->b1.as<float>(), H);
    k_to_f32<bf16><<<64, 256>>>(static_cast<const bf16*>(weights[2]), c->nw.as<float>(), H);
    k_to_bf16<bf16><<<1024, 256>>>(static_cast<const bf16*>(weights[3]), c->w2.as<bf16>(), H * H);
    k_to_f32<bf16><<<64, 256>>>(static_cast<const bf16*>(weights[4]), c->b2.as<float>(), H);
  }
  k_fill_f32<<<64, 256>>>(c->ones.as<float>(), H, 1.0f);
  cudaError_t e = cudaDeviceSynchronize();
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) {
    vv_set_error("vv_conn_create: %s", cudaGetErrorString(e));
    delete c;
    return VV_ERR_CUDA;
  }
  *out = c;
  return VV_OK;
}

void vv_conn_destroy(VVConn* c) { delete c; }

// out[M, out_dim] (bf16) = fc2(norm(fc1(x))) (+ addend[M, out_dim] bf16 when given).  x: [M, in_dim] in io_dtype.
int vv_conn_forward(VVConn* c, const void* x, int M, int io_dtype, const void* addend, void* out, void* stream) {
  VV_REQUIRE(c && x && out, "vv_conn_forward: null argument");
  VV_REQUIRE(M >= 1 && M <= c->max_rows, "vv_conn_forward: M=%d exceeds max_rows=%d", M, c->max_rows);
  VV_REQUIRE(io_dtype == VV_DTYPE_BF16 || io_dtype == VV_DTYPE_F32, "vv_conn_forward: io_dtype");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int H = c->out_dim, I = c->in_dim;
  const bf16* xin = static_cast<const bf16*>(x);
  if (io_dtype == VV_DTYPE_F32) {
    const size_t n = static_cast<size_t>(M) * I;
    VV_CHECK_CUDA(vv_launch(k_to_bf16<float>, g1(n, 256), dim3(256), 0, st, 1, true, static_cast<const float*>(x),
                            c->xin.as<bf16>(), n));
    xin = c->xin.as<bf16>();
  }
  GemmArgs a1;
  a1.A = xin; a1.lda = I; a1.B = c->w1.as<bf16>(); a1.ldb = I; a1.M = M; a1.N = H; a1.K = I;
  a1.mode = GM_STORE_F32; a1.out = c->h.p; a1.ldo = H; a1.bias = c->b1.as<float>();
  a1.b_streaming = 1;
  gemm_plan(a1, vv_num_sms());
  VV_TRY(gemm_launch(a1, st));
  VV_CHECK_CUDA(vv_launch(k_conn_norm, dim3(M), dim3(256), 0, st, 1, true, static_cast<const float*>(c->h.as<float>()),
                          static_cast<const float*>(c->nw.as<float>()), c->y.as<bf16>(), H, c->eps));
  GemmArgs a2;
  a2.A = c->y.as<bf16>(); a2.lda = H; a2.B = c->w2.as<bf16>(); a2.ldb = H; a2.M = M; a2.N = H; a2.K = H;
  a2.out = out; a2.ldo = H; a2.bias = c->b2.as<float>();
  a2.b_streaming = 1;
  if (addend != nullptr) {
    a2.mode = GM_RESID_GAMMA_BF16; a2.gamma = c->ones.as<float>(); a2.resid = addend; a2.ldr = H;
  } else {
    a2.mode = GM_STORE_BF16;
  }
  gemm_plan(a2, vv_num_sms());
  VV_TRY(gemm_launch(a2, st));
  return VV_OK;
}

int vv_latent_sample(const void* mean, const void* noise, const void* row_noise, float std, float bias, float scale,
                     int n, int T, int dim, int io_dtype, void* out, void* std_out, void* stream) {
  VV_REQUIRE(mean && out, "vv_latent_sample: null argument");
  VV_REQUIRE(n >= 1 && T >= 1 && dim >= 1, "vv_latent_sample: n=%d T=%d dim=%d", n, T, dim);
  VV_REQUIRE(io_dtype == VV_DTYPE_BF16 || io_dtype == VV_DTYPE_F32, "vv_latent_sample: io_dtype");
  VV_REQUIRE(noise != nullptr || row_noise == nullptr, "vv_latent_sample: row_noise without noise");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t per_row = static_cast<size_t>(T) * dim, total = per_row * n;
  const unsigned grid = static_cast<unsigned>(std::min<size_t>((total + 255) / 256, 148u * 8u));
  if (io_dtype == VV_DTYPE_BF16)
    VV_CHECK_CUDA(vv_launch(k_latent_sample<bf16>, dim3(grid), dim3(256), 0, st, 1, true, static_cast<const bf16*>(mean),
                            static_cast<const bf16*>(noise), static_cast<const bf16*>(row_noise), std, bias, scale, per_row,
                            total, static_cast<bf16*>(out), static_cast<bf16*>(std_out)));
  else
    VV_CHECK_CUDA(vv_launch(k_latent_sample<float>, dim3(grid), dim3(256), 0, st, 1, true, static_cast<const float*>(mean),
                            static_cast<const float*>(noise), static_cast<const float*>(row_noise), std, bias, scale,
                            per_row, total, static_cast<float*>(out), static_cast<float*>(std_out)));
  return VV_OK;
}

}  // extern "C"
