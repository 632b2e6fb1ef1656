// Issue-rate microbenchmark for the instruction classes the fused decoder kernels are made of (sm_100a).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/pipe_rates scripts/ubench/pipe_rates.cu && gpurun_out/pipe_rates
// One CTA per SM, W warps per scheduler (SMSP), 8 independent dependency chains per thread, N iterations;
// prints clocks per warp-instruction per SMSP.
#include <cstdio>
#include <cuda_runtime.h>
#include <cstdint>

template <int OP>
__global__ void k(float* out, long long* clk, int iters, float seed) {
  float2 a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = make_float2(seed + i + threadIdx.x, seed - i);
  float2 b = make_float2(seed * 0.5f, seed * 0.25f), c = make_float2(seed * 0.125f, 1.0f);
  uint32_t u[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) u[i] = threadIdx.x * 7 + i;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (OP == 0) a[i] = __ffma2_rn(a[i], b, c);                                   // FFMA2 (3 x 64-bit regs)
      if (OP == 1) { a[i].x = fmaf(a[i].x, b.x, c.x); }                             // FFMA 3-reg
      if (OP == 2) { a[i].x = fmaf(a[i].x, 1.0001f, 0.5f); }                        // FFMA imm
      if (OP == 3) a[i] = __fmul2_rn(a[i], b);                                      // FMUL2
      if (OP == 4) { asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(a[i].x)); }  // MUFU.RCP
      if (OP == 5) { asm volatile("prmt.b32 %0, %0, %1, 0x1044;" : "+r"(u[i]) : "r"(u[(i + 1) & 7])); }   // PRMT (alu)
      if (OP == 6) { u[i] = u[i] * 0x10001u + u[(i + 1) & 7]; }                     // IMAD
      if (OP == 7) { a[i].x = fmaxf(a[i].x, b.x) ; a[i].x += 1e-3f; }               // FMNMX + FADD
      if (OP == 8) { a[i] = __ffma2_rn(a[i], b, c); a[i].y = fmaf(a[i].y, b.x, c.x); }   // FFMA2 + FFMA interleaved
      if (OP == 9) { a[i] = __ffma2_rn(a[i], b, c); asm volatile("prmt.b32 %0, %0, %1, 0x1044;" : "+r"(u[i]) : "r"(u[(i + 1) & 7])); }  // FFMA2 + PRMT
      if (OP == 10) { a[i] = __ffma2_rn(a[i], make_float2(1.0001f, 1.0001f), make_float2(0.5f, 0.5f)); }   // FFMA2 with constants
    }
  }
  const long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y + __uint_as_float(u[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, int instr_per_slot, int warps_per_smsp) {
  float* out; long long* clk;
  const int threads = 128 * warps_per_smsp, iters = 2000;
  cudaMalloc(&out, 148 * threads * sizeof(float)); cudaMalloc(&clk, 148 * sizeof(long long));
  k<OP><<<148, threads>>>(out, clk, iters, 1.0f);
  cudaDeviceSynchronize();
  k<OP><<<148, threads>>>(out, clk, iters, 1.0f);
  long long h[148]; cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
  double c = h[74];
  const double n = double(iters) * 8 * instr_per_slot * warps_per_smsp;   // warp-instructions per SMSP
  printf("%-28s warps/SMSP %d: %.2f clk per warp-instruction per SMSP (%s)\n", name, warps_per_smsp, c / n, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out); cudaFree(clk);
}

int main() {
  for (int w : {1, 4}) {
    run<0>("FFMA2 (3-reg)", 1, w);
    run<10>("FFMA2 (const operands)", 1, w);
    run<1>("FFMA (3-reg)", 1, w);
    run<2>("FFMA (imm)", 1, w);
    run<3>("FMUL2", 1, w);
    run<4>("MUFU.RCP", 1, w);
    run<5>("PRMT", 1, w);
    run<6>("IMAD", 1, w);
    run<7>("FMNMX+FADD", 2, w);
    run<8>("FFMA2+FFMA pair", 2, w);
    run<9>("FFMA2+PRMT pair", 2, w);
  }
  return 0;
}
