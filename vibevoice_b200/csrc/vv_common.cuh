// vv_common.cuh — shared host/device helpers for libvvb200 (sm_100a only).
//
// Everything in this library is written for B200 (sm_100a): TMA
// (cp.async.bulk.tensor), mbarrier pipelines, tcgen05.mma with TMEM
// accumulators.  There is no fallback path: if a CUDA call fails the C-ABI
// entry point returns a negative code and vv_last_error() holds the reason.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

typedef __nv_bfloat16 bf16;

// ----------------------------------------------------------------------------
// Error plumbing (host)
// ----------------------------------------------------------------------------
enum VVStatus : int {
  VV_OK = 0,
  VV_ERR_INVALID = -1,   // bad argument (shape / alignment / null pointer)
  VV_ERR_CUDA = -2,      // CUDA runtime / driver failure
  VV_ERR_UNSUPPORTED = -3,
  VV_ERR_STATE = -4,     // call order (e.g. sample before set_schedule)
};

void vv_set_error(const char* fmt, ...);
int vv_num_sms();
bool vv_option_fused_block();
bool vv_option_block_pipe();        // vv_set_option("block_pipe", 0|1): pipelined fused Block1D (vv_blockp.cu) instead of k_block_fused
bool vv_option_gemm_persistent();  // vv_set_option("gemm_persistent", 0|1): persistent kernel for the dense GEMM launches
int vv_option_chain_stages();     // vv_set_option("chain_stages", k): depth of the persistent kernel's activation ring (32 KB stages = K 128; programs built later)
int vv_option_gemm_stream_bn32(); // vv_set_option("gemm_stream_bn32", 0|1): 32-wide tiles for weight-streaming GEMMs with few tiles
int vv_option_gemm_swap();        // vv_set_option("gemm_swap", 0|1): swapped operands for split-K launches with M <= 64
int vv_option_stage_chain();      // vv_set_option("stage_chain", k): wide stages at one step per utterance run as one persistent launch when rows >= k (0: never)
int vv_option_stage_tiled();      // vv_set_option("stage_tiled", rows): C >= 512 stages with several steps per utterance run as one persistent launch when n * steps <= rows (0: never)
int vv_option_dw_rows();          // vv_set_option("dw_rows", r): time steps per CTA of the mixer kernel k_block_dw (0: 8192 / C)
int vv_option_chain_push_barrier();   // vv_set_option("chain_push_barrier", 1|2): cross-proxy fences of k_chain's grid barrier (2: limited to global memory; 1: unqualified)
int vv_option_head_tail_split();  // vv_set_option("head_tail_split", 0|1): the fused step-boundary phase shares every row pair between two CTAs (column halves)
int vv_option_chain_part_rowmajor();  // vv_set_option("chain_part_rowmajor", 0|1): the head program's split-K partials are stored [M][splits][H]
int vv_option_head_chain_rows();   // vv_set_option("head_chain_rows", r): vv_head_sample uses the persistent kernel up to r rows (2 x utterances)
int vv_option_head_fuse_tail();   // vv_set_option("head_fuse_tail", 0|1): the solver-step boundary of the head program as one row-pair phase (1) or as five phases (0)
int vv_option_dw_bulk();          // vv_set_option("dw_bulk", 0|1): the wide stages' mixer kernel stages its input tile with bulk copies (k_block_dw_bulk)
int vv_option_head_conv_walk();   // vv_set_option("head_conv_walk", 0|1): register-blocked output conv (k_head_conv_walk) instead of the staged one
int vv_option_enc_splitk();       // vv_set_option("enc_splitk", 0|1): split-K for the encoder's long-K small-M GEMMs
int vv_option_chain_krot();       // vv_set_option("chain_krot", r): work item on N tile j starts at k-block (j * r) % nkb (0: all start at 0)
int vv_option_chain_keep_layers(); // vv_set_option("chain_keep_layers", k): the head's first k layers' weights are loaded evict-last (kept in L2 across solver steps)
int vv_option_chain_wstages();    // vv_set_option("chain_wstages", k): cap on the depth of its weight ring (16 KB stages; default: what fits)
int vv_option_head_chain();     // vv_set_option("head_chain", 0|1): 0 per-launch path, 1 (or 2) the persistent kernel   // vv_set_option("fused_block", 0|1)
#include <atomic>
extern std::atomic<long long> g_vv_launches;
#define VV_COUNT_LAUNCH() (g_vv_launches.fetch_add(1, std::memory_order_relaxed))
// host synchronisations the LIBRARY initiates (program builds, schedule upload, program-cache eviction): vv_host_sync_count()
extern std::atomic<long long> g_vv_host_syncs;
#define VV_COUNT_SYNC() (g_vv_host_syncs.fetch_add(1, std::memory_order_relaxed))

#define VV_CHECK_CUDA(expr)                                                        \
  do {                                                                             \
    cudaError_t _e = (expr);                                                       \
    if (_e != cudaSuccess) {                                                       \
      vv_set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr,              \
                   cudaGetErrorString(_e));                                        \
      return VV_ERR_CUDA;                                                          \
    }                                                                              \
  } while (0)

#define VV_REQUIRE(cond, ...)                                                      \
  do {                                                                             \
    if (!(cond)) {                                                                 \
      vv_set_error(__VA_ARGS__);                                                   \
      return VV_ERR_INVALID;                                                       \
    }                                                                              \
  } while (0)

#define VV_TRY(expr)                                                               \
  do {                                                                             \
    int _s = (expr);                                                               \
    if (_s != VV_OK) return _s;                                                    \
  } while (0)

static inline int vv_ceil_div(int a, int b) { return (a + b - 1) / b; }

// cudaFuncSetAttribute and the SM count are per DEVICE: one flag / value per device ordinal, looked up with the
// current device (a process may hold handles on several GPUs, e.g. a device_map-split model)
constexpr int VV_MAX_DEVICES = 64;
static inline int vv_current_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= VV_MAX_DEVICES) dev = 0;
  return dev;
}

// Simple owning device buffer used by the handles.
struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  int alloc(size_t n) {
    free_();
    if (n == 0) n = 16;
    cudaError_t e = cudaMalloc(&p, n);
    if (e != cudaSuccess) {
      p = nullptr;
      vv_set_error("cudaMalloc(%zu) failed: %s", n, cudaGetErrorString(e));
      return VV_ERR_CUDA;
    }
    bytes = n;
    return VV_OK;
  }
  int ensure(size_t n) { return (n <= bytes && p) ? VV_OK : alloc(n); }
  void free_() {
    if (p) cudaFree(p);
    p = nullptr;
    bytes = 0;
  }
  template <typename T>
  T* as() const { return reinterpret_cast<T*>(p); }
  ~DevBuf() { free_(); }
  DevBuf() = default;
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
};

// ----------------------------------------------------------------------------
// Device-side PTX wrappers
// ----------------------------------------------------------------------------
#ifdef __CUDACC__

// Launch with programmatic stream serialization (PDL): the kernel may start while its
// predecessor in the stream is still draining; it must call griddep_wait() before touching
// anything the predecessor writes (or anything the predecessor may still be reading, if it
// writes it).  `cluster_x` > 1 adds a cluster dimension.
bool vv_pdl_enabled();
// diagnostic (vv_set_option("launch_log", 1) ... 0): every vv_launch is bracketed by CUDA events; switching it off prints
// kernel name and device time of each launch to stderr (warm caches, but no overlap between launches)
bool vv_launch_log_on();
void vv_launch_log_begin(const void* fn, cudaStream_t st);
void vv_launch_log_end(cudaStream_t st);
template <typename... KArgs, typename... Args>
inline cudaError_t vv_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                             cudaStream_t stream, int cluster_x, bool pdl, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[2];
  unsigned na = 0;
  if (cluster_x > 1) {
    attrs[na].id = cudaLaunchAttributeClusterDimension;
    attrs[na].val.clusterDim.x = static_cast<unsigned>(cluster_x);
    attrs[na].val.clusterDim.y = 1;
    attrs[na].val.clusterDim.z = 1;
    ++na;
  }
  if (pdl && vv_pdl_enabled()) {
    attrs[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attrs;
  cfg.numAttrs = na;
  VV_COUNT_LAUNCH();
  if (vv_launch_log_on()) {
    vv_launch_log_begin(reinterpret_cast<const void*>(kernel), stream);
    const cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
    vv_launch_log_end(stream);
    return e;
  }
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---- mbarrier ---------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a pipeline bug must surface as a trap (launch failure the host
// reports through vv_last_error) and never as a hung GPU box.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 26)) {
      printf("vvb200: mbarrier wait timed out (block %d,%d,%d thread %d)\n", blockIdx.x,
             blockIdx.y, blockIdx.z, threadIdx.x);
      __trap();
    }
  }
}

// Same with a suspend-time hint: the thread may be parked by the hardware for up to `ns` nanoseconds per
// attempt instead of spinning (a polling warp otherwise takes issue slots from the working warps of its
// scheduler: ncu r02, 28 % of k_block_pipe's instructions were try_wait loops).  It still wakes on completion.
__device__ __forceinline__ bool mbar_try_wait_hint(uint64_t* bar, uint32_t parity, uint32_t ns) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity), "r"(ns)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait_parked(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait_hint(bar, parity, 4000u)) {
    if (++spins > (1u << 22)) {
      printf("vvb200: mbarrier wait timed out (block %d,%d,%d thread %d)\n", blockIdx.x, blockIdx.y, blockIdx.z,
             threadIdx.x);
      __trap();
    }
  }
}

// ---- TMA --------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
// L2 cache-policy constants (createpolicy fractional encodings).
#define VV_L2_EVICT_NORMAL 0x1000000000000000ull
#define VV_L2_EVICT_FIRST 0x12F0000000000000ull
#define VV_L2_EVICT_LAST 0x14F0000000000000ull

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* m, uint64_t* bar,
                                            int c0, int c1, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}

// L2 prefetch of one box (no shared memory, no barrier: fire and forget)
__device__ __forceinline__ void tma_prefetch_l2_2d(const CUtensorMap* m, int c0, int c1) {
  asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(reinterpret_cast<uint64_t>(m)), "r"(c0),
               "r"(c1)
               : "memory");
}

// Multicast variant: the box lands at the same CTA-relative offset in every CTA of `mask`
// and signals the mbarrier at the same offset in each of them.
__device__ __forceinline__ void tma_load_2d_mc(void* smem_dst, const CUtensorMap* m, uint64_t* bar,
                                               int c0, int c1, uint16_t mask, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
      ".L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5, %6;" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask), "l"(policy)
      : "memory");
}

// ---- thread-block clusters ---------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// ---- programmatic dependent launch (PDL) --------------------------------------
// wait: every memory operation of the prerequisite grids is complete and visible.
// launch_dependents: the next kernel in the stream may start its prologue (it blocks in its own
// griddepcontrol.wait before touching anything this grid writes).
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ---- tcgen05 / TMEM ---------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; bf16 inputs, fp32 accumulate.
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Warp-converged forms: ALL 32 lanes of the issuing warp execute these; elect.sync picks the one that issues.
// ptxas lowers this pattern to a bare UTCHMMA / UTCBAR.  The single-thread forms above, called from divergent
// code (`if (lane == 0)`), are wrapped by ptxas in an ELECT ... BRA.U.ANY loop per instruction, which costs the
// issuing thread ~68 clk per MMA and ~116 clk per commit (scripts/probe_commit.py, profiles/r01_mma_issue.md) -
// more than the tensor time of a 128 x 64 x 16 MMA (32 clk).
__device__ __forceinline__ void umma_bf16_warp(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                               uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_warp(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void umma_commit_mc_warp(uint64_t* bar, uint16_t mask) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n\t}" ::"r"(
          smem_u32(bar)),
      "h"(mask)
      : "memory");
}
// Arrive on an mbarrier once all previously issued tcgen05.mma have completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
// Same, arriving on the barrier at this offset in every CTA of `mask` (frees a multicast slot).
__device__ __forceinline__ void umma_commit_mc(uint64_t* bar, uint16_t mask) {
  asm volatile(
      "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"(mask)
      : "memory");
}
// 32 lanes x 32 columns of fp32: thread i of the warp receives lane (base+i),
// columns [col, col+32).
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
        "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]),
        "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major, 128-byte-swizzled shared-memory operand descriptor (sm_100 format):
//   [0,14)  start address >> 4        [16,30) leading byte offset >> 4 (unused for SW128 K-major)
//   [32,46) stride byte offset >> 4   [46,48) descriptor version = 1
//   [61,64) layout type (2 = SWIZZLE_128B)
// Rows are 128 B (64 bf16); 8-row groups are 1024 B apart (SBO).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>(1) << 16;
  d |= static_cast<uint64_t>(1024 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}
// kind::f16 instruction descriptor: D=f32, A=B=bf16, both K-major.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(N >> 3) << 17) |
         (static_cast<uint32_t>(M >> 4) << 24);
}

// ---- small math helpers -----------------------------------------------------
// x * sigmoid(x) with the approximate reciprocal (MUFU.RCP, ~1 ulp): the IEEE division's slow path made the SwiGLU
// epilogue of k_chain cost 140 clk per element per thread (2.3 k clk per gate/up phase, scripts/trace_chain.py);
// the result is rounded to bf16 right after.  1 + exp(-x) = inf for x < -88 gives -0, as the exact form does.
__device__ __forceinline__ float silu_f(float x) { return __fdividef(x, 1.0f + __expf(-x)); }
// Exact (erf) GELU as used by ACT2FN["gelu"] in the reference FFN
// (modular_vibevoice_tokenizer.py:589).
__device__ __forceinline__ float gelu_erf_f(float x) {
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f));
}
// Same function through Abramowitz-Stegun 7.1.26 (|erf error| <= 1.5e-7, GELU abs error <= 2.2e-7,
// i.e. < 0.2 ulp of the bf16 value it is rounded to right after, for |x| < 4): one rcp + one ex2
// and ~13 FMA-class instructions, no divergent branches - about half the issue slots of erff().
// The narrow decoder stages are bound by exactly these instructions (26 M GELUs per block).
__device__ __forceinline__ float gelu_erf_fast(float x) {
  const float z = x * 0.70710678118654752f;
  const float az = fabsf(z);
  float t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, az, 1.0f)));
  float poly = fmaf(t, 1.061405429f, -1.453152027f);
  poly = fmaf(poly, t, 1.421413741f);
  poly = fmaf(poly, t, -0.284496736f);
  poly = fmaf(poly, t, 0.254829592f);
  poly *= t;
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(az * az * -1.4426950408889634f));
  const float erf_abs = fmaf(-poly, e, 1.0f);
  const float erf_z = copysignf(erf_abs, z);
  const float hx = 0.5f * x;
  return fmaf(hx, erf_z, hx);
}
// Two GELUs at once on the packed-fp32 pipe (FFMA2 / FMUL2, sm_100).  The fused decoder kernels'
// GELU epilogues are bound by the special-function unit (16 lanes/clk/SM) before anything else, so
// this uses the Abramowitz-Stegun form with ONE MUFU per element (7.1.28):
//   erf(z) = 1 - 1 / (1 + a1 z + ... + a6 z^6)^16,   |err| <= 3e-7, z >= 0
// and x*erf(x/sqrt2) being even removes all sign handling:
//   gelu(x) = max(x, 0) - |x|/2 * (1 + sum_k a_k (|x|/sqrt2)^k)^-16
// |gelu_fast - gelu_exact| <= 7e-7 in fp32 arithmetic over [-12, 12] (scripts check in DESIGN.md section 4);
// the polynomial overflows to +inf for |x| > ~25 and the reciprocal then returns 0 (gelu = relu).
// 18 instructions per PAIR: 2 LOP, 2 FMNMX, 2 MUFU, 12 packed.
__device__ __forceinline__ float2 gelu_erf_fast2(float2 x) {
  float2 u;                                               // -|x|
  u.x = __uint_as_float(__float_as_uint(x.x) | 0x80000000u);
  u.y = __uint_as_float(__float_as_uint(x.y) | 0x80000000u);
#define VV_F2(c) make_float2(c, c)
  float2 P = __ffma2_rn(u, VV_F2(5.3829750000e-06f), VV_F2(-4.8890635643e-05f));   // c_k = a_k * (-1/sqrt2)^k
  P = __ffma2_rn(P, u, VV_F2(3.8003575000e-05f));
  P = __ffma2_rn(P, u, VV_F2(-3.2776263241e-03f));
  P = __ffma2_rn(P, u, VV_F2(2.1141006150e-02f));
  P = __ffma2_rn(P, u, VV_F2(-4.9867346967e-02f));
  P = __ffma2_rn(P, u, VV_F2(1.0f));
#undef VV_F2
  P = __fmul2_rn(P, P);
  P = __fmul2_rn(P, P);
  P = __fmul2_rn(P, P);
  P = __fmul2_rn(P, P);
  float2 r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r.x) : "f"(P.x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r.y) : "f"(P.y));
  const float2 hu = __fmul2_rn(u, make_float2(0.5f, 0.5f));
  return __ffma2_rn(hu, r, make_float2(fmaxf(x.x, 0.f), fmaxf(x.y, 0.f)));
}
// v[i] = gelu(acc[i] + bias[i]) for 8 accumulator words
__device__ __forceinline__ void gelu8_bias(const uint32_t* r, const float (&bv)[8], float (&v)[8]) {
#pragma unroll
  for (int i = 0; i < 8; i += 2) {
    const float2 o = gelu_erf_fast2(__fadd2_rn(make_float2(__uint_as_float(r[i]), __uint_as_float(r[i + 1])),
                                               make_float2(bv[i], bv[i + 1])));
    v[i] = o.x;
    v[i + 1] = o.y;
  }
}
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ void unpack_bf16x8(const uint4& v, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack_bf16x8(const float (&f)[8]) {
  uint4 v;
  v.x = pack_bf16x2(f[0], f[1]);
  v.y = pack_bf16x2(f[2], f[3]);
  v.z = pack_bf16x2(f[4], f[5]);
  v.w = pack_bf16x2(f[6], f[7]);
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
#endif  // __CUDACC__
