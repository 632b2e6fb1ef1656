// vv_chain.cu — persistent phase engine (see vv_chain.cuh).
#include "vv_chain.cuh"

#include <algorithm>
#include <cstring>
#include <vector>

#include "../../include/vvb200.h"
#include "vv_gemm_dev.cuh"

namespace {

constexpr int CH_BN = 64;
constexpr int CH_MAX_A_STAGES = 4;                  // activation ring (also the row phases' staging area)
constexpr int CH_MAX_W_STAGES = 12;                 // weight ring
// A ring stage holds CH_KPS k-blocks (K = 128): waiting on an mbarrier costs the MMA warp ~130-200 clk and a
// tcgen05.commit ~50-100 (profiles/r01_mma_issue.md) against ~47 clk per 128 x 64 x 16 MMA, so the
// synchronisation is paid once per eight MMAs instead of once per four.
#ifndef VV_CH_KPS
#define VV_CH_KPS 2
#endif
constexpr int CH_KPS = VV_CH_KPS;
constexpr int CH_A_TILE = BLOCK_M * BLOCK_K * 2;    // one TMA box of activations
constexpr int CH_B_TILE = CH_BN * BLOCK_K * 2;      // one TMA box of weights
constexpr int CH_A_BYTES = CH_KPS * CH_A_TILE;
constexpr int CH_B_BYTES = CH_KPS * CH_B_TILE;
constexpr int CH_TMEM_COLS = 4 * CH_BN;             // two accumulator buffers x two MMA issuers
constexpr int CH_EPI_THREADS = 32 * NUM_EPI_WARPS;
// warps: 0 activation-tile producer, 1 MMA issuer (even ring groups), 2..9 epilogue / row phases, 10 weight-tile
// producer, 11 MMA issuer (odd ring groups).
// Two issuers: a 128 x 64 x 16 MMA occupies the tensor pipe ~48 clk, but one warp cannot issue a K = 128 group faster
// than ~780-860 clk (two mbarrier waits ~150 each, eight MMAs, two commits ~150: a thread's MMA issue does not overlap
// its own waits and commits), so the main loops ran at half the pipe rate.  Two warps that take alternate ring groups
// into accumulators of their own reach it (scripts/probe_commit.py: 8 MMAs + 2 waits + 2 commits per iteration cost
// EACH of two warps 824 clk, the same as one warp alone); the epilogue adds the two accumulators.
// The weight producer takes no part in the grid barriers (weights depend on nothing computed here).
// Warp order: 0 activation producer, 1 MMA issuer, 2 weight producer, 3 second MMA issuer, 4..11 epilogue / row
// phases - the four light roles form one warpgroup and the eight heavy ones two, so that the register file can be
// re-split between them (setmaxnreg, per warpgroup): 56 registers for the producers and issuers, 224 for the
// epilogue / row-phase warps (128 x 56 + 256 x 224 = 64 512 of 65 536).  At the uniform 168 the row phases could not
// keep a row's loads in flight without spilling.
constexpr int CH_W_WARP = 2;
constexpr int CH_MMA2_WARP = 3;
constexpr int CH_EPI_WARP0 = 4;
constexpr int CH_THREADS = NUM_THREADS + 64;
constexpr int CH_SYNC_THREADS = NUM_THREADS + 32;   // everyone but the weight producer
constexpr int CH_REGS_LIGHT = 96, CH_REGS_HEAVY = 200;
constexpr int CH_CTRL_BYTES = 1024;

#ifdef VV_GEMM_TRACE
// [cta][16*phase + k] for the first 24 phases (k = 8..15: stamps inside row phases, thread 64).  k: 0 epilogue thread 64 finished the phase's work,
// 1 it passed the CTA barrier that opens the grid barrier, 2 warp 1 stored the flag, 3 warp 1 saw all
// flags, 4 thread 64 left the grid barrier, 5 producer lane 0 reached the barrier, 6 MMA lane 0 reached
// it, 7 %globaltimer at 4.
__device__ long long g_chain_trace[160][384];
#define CTRACE(slot, val) do { if ((slot) < 384) g_chain_trace[blockIdx.x][slot] = (val); } while (0)
// per ring group of the first 24 phases: [cta][phase][0: weights landed, 1: activations landed, 2: weight load issued][group < 16]
__device__ long long g_chain_trace2[160][24][3][16];
#define CTRACE2(ph, what, j, val) do { if ((ph) < 24 && (j) < 16) g_chain_trace2[blockIdx.x][ph][what][j] = (val); } while (0)
__device__ __forceinline__ long long gtimer() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#else
#define CTRACE(slot, val) do { } while (0)
#define CTRACE2(ph, what, j, val) do { } while (0)
#endif

enum ChainOp : int { CH_GEMM = 0, CH_NORM_MOD = 1, CH_CFG = 2, CH_TAIL = 3, CH_DWMIX = 4, CH_DWTILE = 5 };

struct alignas(16) ChainPhase {   // one entry of the program (copied to shared memory at kernel start)
  int op;
  // CH_GEMM
  int tmA, tmB;              // indices into the tensor-map table
  int n_tiles_n, n_items;    // N tiles; work items = n_tiles_n * n_tiles_m * splits (item = (z * n_tiles_m + m) * n_tiles_n + n)
  int n_tiles_m;             // 128-row tiles of A / the output
  int part_zs;               // GM_PARTIAL_F32: elements between two splits' outputs of a row (0: splits are whole [M][ldo] planes)
  union {
    GemmKParams g;           // CH_GEMM
    ChainNormMod nm;         // CH_NORM_MOD
    ChainCfg cf;             // CH_CFG
    ChainTail tl;            // CH_TAIL
    ChainDw dw;              // CH_DWMIX
    ChainDwTile dt;          // CH_DWTILE
  };
};
static_assert(sizeof(ChainPhase) % 16 == 0 && sizeof(ChainPhase) <= 224, "program entry size");

// What the producers and the MMA issuers need of a phase (16 bytes).  The full entries no longer fit in shared memory
// next to a 4-deep activation ring and a 6-deep weight ring (122 entries = 23 KB for the head), so these three roles
// read this table from global memory - each fetches the NEXT phase's entry before it enters the grid barrier - and
// only the epilogue warps see full entries, through a two-slot window filled by bulk copies one phase ahead.
struct alignas(16) ChainLite {
  uint8_t op;
  uint8_t flag;              // CH_GEMM: weight-tile L2 policy (0 evict-first, 1 evict-last); CH_TAIL: `last`
  uint8_t n_tiles_m;
  uint8_t a_rows;
  int16_t tmA, tmB;
  uint16_t n_items;          // CH_TAIL: n (row pairs)
  uint16_t n_tiles_n;
  uint16_t kb_total;         // k-blocks of the whole K; CH_TAIL: H / 64
  uint16_t kb_per_split;
};
static_assert(sizeof(ChainLite) == 16, "lite entry");
// volatile: the load must be issued where it is written (before the grid barrier); ptxas otherwise sinks a plain
// read-only load to its first use behind the barrier, an L2 round trip on the critical path (+580 clk per phase, traced)
__device__ __forceinline__ ChainLite lite_ld(const ChainLite* t, int i) {
  uint4 v;
  asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(t + i));
  return *reinterpret_cast<const ChainLite*>(&v);
}
// the same entry for every lane, in a form ptxas can prove warp-uniform (an MMA issuer's loop bounds must be)
__device__ __forceinline__ ChainLite lite_uniform(const ChainLite& l) {
  uint4 v = *reinterpret_cast<const uint4*>(&l);
  v.x = __shfl_sync(0xffffffffu, v.x, 0); v.y = __shfl_sync(0xffffffffu, v.y, 0);
  v.z = __shfl_sync(0xffffffffu, v.z, 0); v.w = __shfl_sync(0xffffffffu, v.w, 0);
  return *reinterpret_cast<const ChainLite*>(&v);
}

struct ChainSmem {           // control block behind the rings
  uint64_t full_a[CH_MAX_A_STAGES];
  uint64_t empty_a[CH_MAX_A_STAGES];
  uint64_t full_b[CH_MAX_W_STAGES];
  uint64_t empty_b[CH_MAX_W_STAGES];
  uint64_t acc_full[2];
  uint64_t acc_empty[2];
  uint64_t row_bar[5];       // CH_ROW_BARS: bulk-copy staging of the row phases
  uint64_t win_bar[2];       // the epilogue warps' window of full program entries
  uint32_t tmem_slot;
  float red[2 * NUM_EPI_WARPS];
  alignas(16) ChainPhase win[2];
};
static_assert(sizeof(ChainSmem) <= CH_CTRL_BYTES, "control block");
constexpr int CH_MIN_A_STAGES = 2;
constexpr int CH_MIN_W_STAGES = 2;
// dynamic smem = 1024 (alignment) + SA * A_BYTES + SB * B_BYTES + CTRL_BYTES + program
__device__ __forceinline__ void cta_sync_no_w() { asm volatile("bar.sync 2, %0;" ::"n"(CH_SYNC_THREADS) : "memory"); }

__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }

// Grid barrier between phases.  Every thread of every CTA but the weight producer calls it with the same epoch.
// Strong (gpu-scope) loads of ONE line by all 148 CTAs serialise at its L2 slice: a counter that everyone spins on,
// or a flag array that everyone scans, both measured 6-13k clk per barrier (profiles/r01_chain_trace.md).  A counter
// plus a flag broadcast by the last arriver is three dependent L2 trips (atomic add and its return value -> flag
// stores -> the waiter's poll; 1.9 us per barrier with skew).  So: every arriving CTA adds 1 to a private counter line
// of EVERY CTA (fire-and-forget reductions, one warp, ~5 per lane) and then polls only its own line until it holds
// epoch * gridDim - one trip for the reductions plus the poll that observes them (head 1.113 -> 1.038 ms).  The
// counters are never reset: they count all barriers of the program's life, modulo 2^32.
// Ordering: data stores -> CTA barrier -> cross-proxy fence + gpu-scope fence by the leader warp, then its reductions
// (cumulative release, the pattern cooperative-groups' grid.sync uses) -> waiter: acquire load of its own counter ->
// cross-proxy fence (the next phase reads through TMA) -> CTA barrier.  The cross-proxy fences are limited to the
// global state space: an unqualified fence.proxy.async also waits on shared memory and cost ~0.4 us per barrier
// (head 1.044 -> 0.995 ms); staging in shared memory that a TMA overwrites later has its own fence_proxy_async_smem
// where it is written.  fences == 1 keeps the unqualified form (vv_set_option("chain_push_barrier", 1), for A/B runs).
// The spin is bounded: a scheduling bug must surface as a trap, never as a hung GPU.
// The LEADER is the activation-tile producer warp: it leaves through bar.arrive, without waiting for the other warps
// to wake up, and issues the next phase's first activation tiles at once (their L2 latency, ~2 k clk, is on every
// GEMM phase's critical path).  Entry and exit use different named barriers, so the leader's next entry cannot be
// counted into an exit that a slow warp has not reached yet.
constexpr int CH_FLAG_STRIDE = 32;   // uint32 words between two CTAs' counters (128 B)
constexpr int CH_MAX_GRID = 512;     // counter lines allocated per program
__device__ __forceinline__ void cta_exit_sync() { asm volatile("bar.sync 3, %0;" ::"n"(CH_SYNC_THREADS) : "memory"); }
__device__ __forceinline__ void cta_exit_arrive() { asm volatile("bar.arrive 3, %0;" ::"n"(CH_SYNC_THREADS) : "memory"); }
template <bool LEADER>
__device__ __forceinline__ void grid_sync(uint32_t* sync, uint32_t epoch, int ph, int fences) {
  if (threadIdx.x == 0) CTRACE(16 * ph + 5, clock64());
  if (threadIdx.x == 32) CTRACE(16 * ph + 6, clock64());
  cta_sync_no_w();
  if (threadIdx.x == 32 * CH_EPI_WARP0) CTRACE(16 * ph + 1, clock64());
  if constexpr (LEADER) {
    const int lane = threadIdx.x & 31;
    uint32_t* lines = sync + CH_FLAG_STRIDE;
    if (lane == 0) {
      if (fences == 1) fence_proxy_async_all();
      else fence_proxy_async_global();
    }
    __threadfence();                          // one warp-wide MEMBAR: every lane's reduction below is a release
    for (int c = lane; c < static_cast<int>(gridDim.x); c += 32)
      asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(lines + c * CH_FLAG_STRIDE), "r"(1u) : "memory");
    if (lane == 0) {
      CTRACE(16 * ph + 2, clock64());
      const uint32_t target = epoch * gridDim.x;
      const long long t0 = clock64();
      while (ld_acquire_u32(lines + blockIdx.x * CH_FLAG_STRIDE) != target) {
        if (clock64() - t0 > (1ll << 32)) {      // ~2 s
          printf("vvb200: grid barrier timed out (block %d, epoch %u)\n", blockIdx.x, epoch);
          __trap();
        }
      }
      CTRACE(16 * ph + 3, clock64());
      if (fences == 1) fence_proxy_async_all();
      else fence_proxy_async_global();
    }
    __syncwarp();
    cta_exit_arrive();
  } else {
    cta_exit_sync();
  }
}

// 32 lanes x 16 consecutive fp32 columns of TMEM (tcgen05.ld.32x32b.x16)
__device__ __forceinline__ void tmem_ld_16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                 "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(taddr));
}
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(CH_EPI_THREADS) : "memory"); }

__device__ __forceinline__ float4 ldcg4(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }

// ---- row phases: staging -------------------------------------------------------------------
// The row phases read whole contiguous rows (a residual row, its split-K partials, the adaLN
// shift/scale/gate rows, small weight matrices).  Register-staged LDGs do badly here: in this
// large kernel ptxas keeps 2-3 loads in flight per thread and spills (ncu source view, r01), and a
// row phase then costs a handful of serialised L2 round trips.  So rows are staged with 1-D bulk
// copies (cp.async.bulk, the TMA engine, no tensor map): one thread issues every copy of a row up
// front, all of them complete on one mbarrier, and the 256 threads then work out of shared memory.
// Staging lives in the A half of the TMA ring, which is idle between GEMM phases (only weight tiles
// run ahead of a grid barrier).
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
constexpr int CH_ROW_BARS = 5;      // [0]: row staging; [1..4]: weight-chunk ring of the tail phase
struct RowCtx {                     // per-thread view of the row-phase machinery
  uint64_t* bars;                   // ChainSmem::row_bar
  uint32_t parity;                  // bit j: phase parity of bars[j] (every epilogue thread tracks all of them)
  uint8_t* scr; int scr_bytes;      // scratch/staging region (16-byte aligned)
  float* red;                       // [16]
  int tid;
  int ph;                           // current phase (trace builds)
};
__device__ __forceinline__ void row_wait(RowCtx& c, int j) {
  mbar_wait(&c.bars[j], (c.parity >> j) & 1u);
  c.parity ^= 1u << j;
}
// staging is about to be overwritten by the async proxy: order this thread's generic accesses first
__device__ __forceinline__ void row_release(RowCtx&) {
  fence_proxy_async_smem();
  epi_bar_sync();
}
__device__ __forceinline__ float block_sum256(RowCtx& c, float v, int slot) {
  v = warp_sum(v);
  if ((c.tid & 31) == 0) c.red[slot * 8 + (c.tid >> 5)] = v;
  epi_bar_sync();
  float t = 0.f;
#pragma unroll
  for (int i = 0; i < NUM_EPI_WARPS; ++i) t += c.red[slot * 8 + i];
  return t;
}
// bf16x4( x * rms * w * (1 + scale) + shift )
__device__ __forceinline__ uint2 modulate4(const float4 x4, float rms, const float4 wv, const uint2 sh, const uint2 sc) {
  const float2 sh0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&sh.x));
  const float2 sh1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&sh.y));
  const float2 sc0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&sc.x));
  const float2 sc1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&sc.y));
  uint2 o;
  o.x = pack_bf16x2(x4.x * rms * wv.x * (1.f + sc0.x) + sh0.x, x4.y * rms * wv.y * (1.f + sc0.y) + sh0.y);
  o.y = pack_bf16x2(x4.z * rms * wv.z * (1.f + sc1.x) + sh1.x, x4.w * rms * wv.w * (1.f + sc1.y) + sh1.y);
  return o;
}

// Stage one residual row and what folds into it, then x += gate * sum_z partial_z in shared memory.
// Layout at `stg`: x [H f32] | shift [H bf16] | scale [H bf16] | gate [H bf16] | partial [nsplit][H f32].
// Returns this thread's share of sum(x^2); the folded row stays in stg (x) and, if xg != null, goes
// back to global memory.
__host__ __device__ inline int row_stage_bytes(int H, int nsplit) { return H * (10 + 4 * nsplit); }
// Layout at `stg`: x [H f32] | gate [H bf16] | shift [H bf16] | scale [H bf16] | partial [nsplit][H f32].
// A bulk copy costs ~250 clk of the copy engine whatever its size (eleven copies held a row back ~3 k clk,
// scripts/trace_chain.py), so a row is fetched with as few as its layout allows: the adaLN table keeps
// [... gate(l-1) | shift(l) | scale(l) ...] of a row side by side (one copy of 3 H), and row-major partials
// (chain_add_gemm's part_rowmajor) are one block of nsplit * H floats.  One copy per warp, lane 0: the lanes of one
// warp would issue theirs one after the other.  The transaction count may run negative until thread 0's expect_tx
// lands: the phase cannot complete before that arrival.
__device__ __forceinline__ void row_issue(RowCtx& c, int bar, uint8_t* stg, const float* x, const float* partial,
                                          int nsplit, int M, int m, const bf16* gate, const bf16* shift,
                                          const bf16* scale, int ld_mod, int H, int rowmajor) {
  if ((c.tid & 31) != 0) return;
  const int w = c.tid >> 5;
  float* sx = reinterpret_cast<float*>(stg);
  bf16* sg = reinterpret_cast<bf16*>(stg + 4 * H);
  float* sp = reinterpret_cast<float*>(stg + 10 * H);
  uint64_t* b = &c.bars[bar];
  if (c.tid == 0) mbar_arrive_expect_tx(b, static_cast<uint32_t>(H * (8 + (nsplit > 0 ? 2 + 4 * nsplit : 0))));
  const bf16* shr = shift + static_cast<size_t>(m) * ld_mod;
  const bf16* scr = scale + static_cast<size_t>(m) * ld_mod;
  const bf16* gr = (nsplit > 0) ? gate + static_cast<size_t>(m) * ld_mod : nullptr;
  if (w == 0) bulk_g2s(sx, x + static_cast<size_t>(m) * H, H * 4, b);
  if (nsplit > 0 && gr + H == shr && shr + H == scr) {
    if (w == 1) bulk_g2s(sg, gr, H * 6, b);
  } else {
    if (w == 1) {
      if (shr + H == scr) {
        bulk_g2s(sg + H, shr, H * 4, b);
      } else {
        bulk_g2s(sg + H, shr, H * 2, b);
        bulk_g2s(sg + 2 * H, scr, H * 2, b);
      }
    }
    if (w == 2 && nsplit > 0) bulk_g2s(sg, gr, H * 2, b);
  }
  if (nsplit > 0) {
    if (rowmajor) {
      if (w == 3) bulk_g2s(sp, partial + static_cast<size_t>(m) * nsplit * H, static_cast<uint32_t>(nsplit) * H * 4, b);
    } else {
      for (int z = (w + NUM_EPI_WARPS - 3) % NUM_EPI_WARPS; z < nsplit; z += NUM_EPI_WARPS)
        bulk_g2s(sp + static_cast<size_t>(z) * H, partial + (static_cast<size_t>(z) * M + m) * H, H * 4, b);
    }
  }
}
__device__ __forceinline__ float row_fold(RowCtx& c, int bar, uint8_t* stg, int nsplit, int m, int H, float* xg) {
  float* sx = reinterpret_cast<float*>(stg);
  const bf16* sg = reinterpret_cast<const bf16*>(stg + 4 * H);
  const float* sp = reinterpret_cast<const float*>(stg + 10 * H);
  row_wait(c, bar);
  float ss = 0.f;
  const int nvec = H / 4;
#pragma unroll 1
  for (int i = c.tid; i < nvec; i += CH_EPI_THREADS) {
    float4 x4 = reinterpret_cast<const float4*>(sx)[i];
    if (nsplit > 0) {
      float4 s4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 2
      for (int z = 0; z < nsplit; ++z) {
        const float4 q = reinterpret_cast<const float4*>(sp + static_cast<size_t>(z) * H)[i];
        s4.x += q.x; s4.y += q.y; s4.z += q.z; s4.w += q.w;
      }
      const uint2 gv = reinterpret_cast<const uint2*>(sg)[i];
      const float2 g0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gv.x));
      const float2 g1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gv.y));
      x4.x += g0.x * s4.x; x4.y += g0.y * s4.y; x4.z += g1.x * s4.z; x4.w += g1.y * s4.w;
      reinterpret_cast<float4*>(sx)[i] = x4;
      if (xg != nullptr) reinterpret_cast<float4*>(xg + static_cast<size_t>(m) * H)[i] = x4;
    }
    ss += x4.x * x4.x + x4.y * x4.y + x4.z * x4.z + x4.w * x4.w;
  }
  return ss;
}

// ---- row phase: split-K reduce + gated residual + RMSNorm + adaLN modulate (k_norm_mod, vv_head.cu) ----
// rows are spread over the CTAs; one row is handled by the 256 epilogue threads of a CTA.
__device__ __forceinline__ void phase_norm_mod(const ChainNormMod& a, RowCtx& c, int ph) {
  const int H = a.H, nvec = H / 4;
  uint8_t* stg = c.scr;
  float* sw = reinterpret_cast<float*>(stg + row_stage_bytes(H, a.nsplit));   // norm weight [H f32]
#pragma unroll 1
  for (int m = blockIdx.x; m < a.M; m += gridDim.x) {
    if (c.tid == 32 * (NUM_EPI_WARPS - 1) && a.w != nullptr) {   // the last warp (the row's copies start at warp 0)
      mbar_arrive_expect_tx(&c.bars[1], H * 4);
      bulk_g2s(sw, a.w, H * 4, &c.bars[1]);
    }
    row_issue(c, 0, stg, a.x, a.partial, a.nsplit, a.M, m, a.prev_gate, a.shift, a.scale, a.ld_mod, H, a.rowmajor);
    const float ss = row_fold(c, 0, stg, a.nsplit, m, H, a.x);
    if (c.tid == 0) CTRACE(16 * ph + 8, clock64());
    const float rms = rsqrtf(block_sum256(c, ss, 0) / static_cast<float>(H) + a.eps);
    if (a.w != nullptr) row_wait(c, 1);
    if (c.tid == 0) CTRACE(16 * ph + 9, clock64());
    const float* sx = reinterpret_cast<const float*>(stg);
    const bf16* ssh = reinterpret_cast<const bf16*>(stg + 6 * H);
    const bf16* ssc = ssh + H;
#pragma unroll 1
    for (int i = c.tid; i < nvec; i += CH_EPI_THREADS) {
      const float4 wv = (a.w != nullptr) ? reinterpret_cast<const float4*>(sw)[i] : make_float4(1.f, 1.f, 1.f, 1.f);
      *reinterpret_cast<uint2*>(a.y + static_cast<size_t>(m) * H + i * 4) =
          modulate4(reinterpret_cast<const float4*>(sx)[i], rms, wv, reinterpret_cast<const uint2*>(ssh)[i],
                    reinterpret_cast<const uint2*>(ssc)[i]);
    }
    if (c.tid == 0) CTRACE(16 * ph + 10, clock64());
    row_release(c);
  }
}

// ---- row phase: Block1D mixer for one new time step per utterance (ChainDw, vv_chain.cuh) ----
// One utterance = one row of C channels = one CTA pass; a thread owns 8 channels.  No shared memory besides the
// reduction scratch: the 7-tap stencil runs over six history rows read from the cache and the row just normalised.
__device__ __forceinline__ void phase_dwmix(const ChainDw& a, const CacheView& cv, RowCtx& c) {
  const int C = a.C, lpr = C >> 3;
  const int cg = c.tid;                       // channel group (C <= 2048: at most one per thread)
  const bool live = cg < lpr;
  const int c0 = cg * 8;
#pragma unroll 1
  for (int b = blockIdx.x; b < a.rows; b += gridDim.x) {
    const size_t o = static_cast<size_t>(b) * C + c0;
    float x[8], xn[8];
    float ss = 0.f;
    if (live) {
      if (a.part == nullptr) {
        unpack_bf16x8(__ldcg(reinterpret_cast<const uint4*>(a.x_in + o)), x);
      } else {
        float s8[8];
        {
          const float4 b0 = __ldg(reinterpret_cast<const float4*>(a.bias2 + c0)), b1 = __ldg(reinterpret_cast<const float4*>(a.bias2 + c0) + 1);
          s8[0] = b0.x; s8[1] = b0.y; s8[2] = b0.z; s8[3] = b0.w; s8[4] = b1.x; s8[5] = b1.y; s8[6] = b1.z; s8[7] = b1.w;
        }
        const size_t ps = static_cast<size_t>(a.rows) * C;
        for (int z = 0; z < a.nsplit; ++z) {
          const float4 p0 = ldcg4(a.part + z * ps + o), p1 = ldcg4(a.part + z * ps + o + 4);
          s8[0] += p0.x; s8[1] += p0.y; s8[2] += p0.z; s8[3] += p0.w; s8[4] += p1.x; s8[5] += p1.y; s8[6] += p1.z; s8[7] += p1.w;
        }
        float xr[8], g2[8];
        unpack_bf16x8(__ldcg(reinterpret_cast<const uint4*>(a.xres + o)), xr);
        {
          const float4 g0 = __ldg(reinterpret_cast<const float4*>(a.gamma2 + c0)), g1 = __ldg(reinterpret_cast<const float4*>(a.gamma2 + c0) + 1);
          g2[0] = g0.x; g2[1] = g0.y; g2[2] = g0.z; g2[3] = g0.w; g2[4] = g1.x; g2[5] = g1.y; g2[6] = g1.z; g2[7] = g1.w;
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) xr[k] += g2[k] * s8[k];
        unpack_bf16x8(pack_bf16x8(xr), x);                 // the residual stream is stored in bf16
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) ss += x[k] * x[k];
    }
    const float rms = rsqrtf(block_sum256(c, ss, 0) / static_cast<float>(C) + a.eps);
    float o8[8];
    float ss2 = 0.f;
    if (live) {
      {
        const float4 w0 = __ldg(reinterpret_cast<const float4*>(a.norm_w + c0)), w1 = __ldg(reinterpret_cast<const float4*>(a.norm_w + c0) + 1);
        const float wn[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
        for (int k = 0; k < 8; ++k) xn[k] = x[k] * rms * wn[k];
      }
      const uint4 xnv = pack_bf16x8(xn);
      unpack_bf16x8(xnv, xn);                              // the stencil reads the stored (bf16) normalised rows
      const bf16* hist = cache_rd(cv, b, a.off);
      bf16* hnew = cache_wr(cv, b, a.off);
      float acc[8];
      {
        const float4 b0 = __ldg(reinterpret_cast<const float4*>(a.dw_b + c0)), b1 = __ldg(reinterpret_cast<const float4*>(a.dw_b + c0) + 1);
        acc[0] = b0.x; acc[1] = b0.y; acc[2] = b0.z; acc[3] = b0.w; acc[4] = b1.x; acc[5] = b1.y; acc[6] = b1.z; acc[7] = b1.w;
      }
#pragma unroll
      for (int j = 0; j <= 6; ++j) {
        float v[8];
        if (j < 6) {
          const uint4 hv = *reinterpret_cast<const uint4*>(hist + static_cast<size_t>(j) * C + c0);
          unpack_bf16x8(hv, v);
          if (j > 0) *reinterpret_cast<uint4*>(hnew + static_cast<size_t>(j - 1) * C + c0) = hv;
        } else {
#pragma unroll
          for (int k = 0; k < 8; ++k) v[k] = xn[k];
          *reinterpret_cast<uint4*>(hnew + static_cast<size_t>(5) * C + c0) = xnv;
        }
        const float4 t0 = __ldg(reinterpret_cast<const float4*>(a.dw_w + static_cast<size_t>(j) * C + c0));
        const float4 t1 = __ldg(reinterpret_cast<const float4*>(a.dw_w + static_cast<size_t>(j) * C + c0) + 1);
        const float wt[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = fmaf(wt[k], v[k], acc[k]);
      }
      {
        const float4 g0 = __ldg(reinterpret_cast<const float4*>(a.gamma + c0)), g1 = __ldg(reinterpret_cast<const float4*>(a.gamma + c0) + 1);
        const float gm[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
#pragma unroll
        for (int k = 0; k < 8; ++k) o8[k] = x[k] + gm[k] * acc[k];
      }
      const uint4 ov = pack_bf16x8(o8);
      *reinterpret_cast<uint4*>(a.x_out + o) = ov;
      unpack_bf16x8(ov, o8);                               // the FFN norm sees the stored (bf16) residual stream
#pragma unroll
      for (int k = 0; k < 8; ++k) ss2 += o8[k] * o8[k];
    }
    const float rms2 = rsqrtf(block_sum256(c, ss2, 1) / static_cast<float>(C) + a.eps);
    if (live) {
      const float4 w0 = __ldg(reinterpret_cast<const float4*>(a.ffn_norm_w + c0)), w1 = __ldg(reinterpret_cast<const float4*>(a.ffn_norm_w + c0) + 1);
      const float wf[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
      for (int k = 0; k < 8; ++k) o8[k] = o8[k] * rms2 * wf[k];
      *reinterpret_cast<uint4*>(a.y + o) = pack_bf16x8(o8);
    }
    epi_bar_sync();                                        // the reduction scratch is reused by the next row
  }
}

// ---- row phase: Block1D mixer over tiles of TT time steps (ChainDwTile, vv_chain.cuh; k_block_dw's arithmetic) ----
__device__ __forceinline__ float rows_sum(RowCtx& c, float v, int lpr, int slot) {
  // sum over the lpr lanes that share a row: sub-warp shuffle, or across warps through shared memory
  if (lpr <= 32) {
    for (int o = lpr >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
  }
  v = warp_sum(v);
  epi_bar_sync();                        // protect the scratch from the previous use
  if ((c.tid & 31) == 0) c.red[slot * 8 + (c.tid >> 5)] = v;
  epi_bar_sync();
  const int wpr = lpr >> 5;              // warps per row
  const int w0 = ((c.tid >> 5) / wpr) * wpr;
  float s = 0.f;
  for (int i = 0; i < wpr; ++i) s += c.red[slot * 8 + w0 + i];
  return s;
}
__device__ __forceinline__ void ldg8(const float* p, float (&f)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
__device__ __forceinline__ void phase_dwtile(const ChainDwTile& a, const CacheView& cv, RowCtx& c) {
  const int C = a.C, Tn = a.Tn, TT = a.TT, tid = c.tid;
  const int lpr = C >> 3;                                  // lanes per row (8 channels each)
  const int lane = tid % lpr, rslot = tid / lpr;
  const int rpp = CH_EPI_THREADS / lpr;                    // rows per pass (>= 1: C <= 2048)
  const int c0 = lane * 8;
  const int tiles = (Tn + TT - 1) / TT;
  bf16* xn = reinterpret_cast<bf16*>(c.scr);               // [(TT + 6)][C]
  float wn[8], bb[8], gm[8], wf[8];
  ldg8(a.norm_w + c0, wn);
  ldg8(a.dw_b + c0, bb);
  ldg8(a.gamma + c0, gm);
  ldg8(a.ffn_norm_w + c0, wf);
#pragma unroll 1
  for (int item = blockIdx.x; item < a.n * tiles; item += gridDim.x) {
    const int b = item / tiles, t0 = (item % tiles) * TT;
    const bf16* xin = a.x_in + static_cast<size_t>(b) * Tn * C;
    const int rows_here = min(TT, Tn - t0);
    // ---- normalised rows (time t0-6 .. t0+rows_here-1) into shared memory ----
    const int nrows = rows_here + HIST;
    const int passes = (nrows + rpp - 1) / rpp;
#pragma unroll 1
    for (int ps = 0; ps < passes; ++ps) {
      const int r = ps * rpp + rslot;
      const int tg = t0 - HIST + r;
      const bool live = r < nrows;
      const bool from_x = live && tg >= 0;
      float f[8];
      float ss = 0.f;
      uint4 hv = make_uint4(0u, 0u, 0u, 0u);
      if (from_x) {
        unpack_bf16x8(__ldcg(reinterpret_cast<const uint4*>(xin + static_cast<size_t>(tg) * C + c0)), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) ss += f[i] * f[i];
      } else if (live) {                                   // rows before this call: already normalised, from the cache
        hv = *reinterpret_cast<const uint4*>(cache_rd(cv, b, a.off) + (HIST + tg) * C + c0);
      }
      ss = rows_sum(c, ss, lpr, 0);
      if (from_x) {
        const float rms = rsqrtf(ss / static_cast<float>(C) + a.eps);
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = f[i] * rms * wn[i];
        hv = pack_bf16x8(f);
      }
      if (live) *reinterpret_cast<uint4*>(xn + static_cast<size_t>(r) * C + c0) = hv;
    }
    epi_bar_sync();
    // ---- new history: normalised rows of time Tn-6 .. Tn-1 (the tile that holds the last row) ----
    if (t0 + TT >= Tn) {
      bf16* dst = cache_wr(cv, b, a.off);
      for (int i = tid; i < HIST * lpr; i += CH_EPI_THREADS) {
        const int q = i / lpr, l = i % lpr;
        *reinterpret_cast<uint4*>(dst + q * C + l * 8) =
            *reinterpret_cast<const uint4*>(xn + static_cast<size_t>(Tn - t0 + q) * C + l * 8);
      }
    }
    // ---- 7-tap depthwise conv, layer-scale residual, FFN norm ----
    const int passes2 = (rows_here + rpp - 1) / rpp;
#pragma unroll 1
    for (int ps = 0; ps < passes2; ++ps) {
      const int r = ps * rpp + rslot;
      const bool live = r < rows_here;
      const int t = t0 + r;
      float o[8];
      float ss = 0.f;
      if (live) {
        float acc[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = bb[i];
#pragma unroll
        for (int j = 0; j <= HIST; ++j) {
          float v[8], wt[8];
          unpack_bf16x8(*reinterpret_cast<const uint4*>(xn + static_cast<size_t>(r + j) * C + c0), v);
          ldg8(a.dw_w + static_cast<size_t>(j) * C + c0, wt);
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[i] = fmaf(wt[i], v[i], acc[i]);
        }
        float xr[8];
        unpack_bf16x8(__ldcg(reinterpret_cast<const uint4*>(xin + static_cast<size_t>(t) * C + c0)), xr);
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = xr[i] + gm[i] * acc[i];
        const uint4 ov = pack_bf16x8(o);
        *reinterpret_cast<uint4*>(a.x_out + (static_cast<size_t>(b) * Tn + t) * C + c0) = ov;
        unpack_bf16x8(ov, o);                 // the FFN norm sees the stored (bf16) residual stream
#pragma unroll
        for (int i = 0; i < 8; ++i) ss += o[i] * o[i];
      }
      ss = rows_sum(c, ss, lpr, 1);
      if (live) {
        const float rms = rsqrtf(ss / static_cast<float>(C) + a.eps);
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = o[i] * rms * wf[i];
        *reinterpret_cast<uint4*>(a.y + (static_cast<size_t>(b) * Tn + t) * C + c0) = pack_bf16x8(o);
      }
    }
    row_release(c);                                        // xn is overwritten by the next item (or by TMA after the phase)
  }
}

// ---- row phase: CFG combine + v -> x0 + DPM-Solver++ update (k_cfg_update, vv_head.cu) ----
__device__ __forceinline__ void phase_cfg(const ChainCfg& a, const ChainCall& call, int tid) {
  const int nL = a.n * a.L;
  const ChainCfg& k = a;
  const size_t MN = static_cast<size_t>(2) * nL;
  for (int i = blockIdx.x * CH_EPI_THREADS + tid; i < nL; i += gridDim.x * CH_EPI_THREADS) {
    const float x = __ldcg(a.zstate + i);
    const float xp0 = (k.order == 2) ? __ldcg(a.prev_x0 + i) : 0.f;
    float c = 0.f, u = 0.f;
    for (int z0 = 0; z0 < a.nsplit; z0 += 8) {          // independent loads first, adds after
      float cq[8], uq[8];
#pragma unroll
      for (int w = 0; w < 8; ++w) {
        const bool on = z0 + w < a.nsplit;
        cq[w] = on ? __ldcg(a.v + (z0 + w) * MN + i) : 0.f;
        uq[w] = on ? __ldcg(a.v + (z0 + w) * MN + nL + i) : 0.f;
      }
#pragma unroll
      for (int w = 0; w < 8; ++w) { c += cq[w]; u += uq[w]; }
    }
    const float veps = u + call.cfg * (c - u);
    const float x0 = k.alpha_s * x - k.sigma_s * veps;
    float xn = k.cx * x + k.c0 * x0;
    if (k.order == 2) xn += k.c1 * (k.inv_r0 * (x0 - xp0));
    if (k.cn != 0.f && call.sde_noise != nullptr) xn += k.cn * call.sde_noise[static_cast<size_t>(a.step) * nL + i];
    a.prev_x0[i] = x0;
    a.zstate[i] = xn;
    const bf16 b = __float2bfloat16(xn);
    a.zb[i] = b;
    a.zb[nL + i] = b;
    if (a.last && call.out != nullptr) {
      if (call.out_dtype == VV_DTYPE_BF16) reinterpret_cast<bf16*>(call.out)[i] = b;
      else reinterpret_cast<float*>(call.out)[i] = xn;
    }
  }
}


// ---- row-pair phase: final layer -> CFG -> solver update -> input projection -> layer-0 norm ----
// One CTA owns the (cond, uncond) rows of one utterance, so nothing in it needs another CTA.  The two
// small Linears (H x 64 each way) have 1-2 activation rows: their weights stream through a ring of
// TMA tiles ([64 x 64] bf16, SWIZZLE_128B - the same tensor maps a GEMM phase would use) and the
// eight warps contract them with warp-level mma.sync.m16n8k16 (weights as the 16-row operand, the
// activation rows as columns 0/1 of the 8-wide operand), so the phase is bound by the weight
// transfer, not by CUDA-core dot products (2 warps per scheduler cannot hide FMA chains: 24k clk
// measured for the scalar version).
// scratch: ybf [2][H] bf16 | xs [H] f32 | vs [2][64] | zs [64] | (pad to 1 KB) | vsw [8 warps][2][64] f32 | staging ...
constexpr int CH_TAIL_FIXED(int H) { return 8 * H + 1024 + 4096; }
constexpr int CH_WTILE = 64 * 64 * 2;            // one [64 x 64] bf16 tile == one weight-ring tile (CH_B_TILE)
__host__ __device__ inline int tail_n0_bytes(int H) { return (12 * H + 1023) / 1024 * 1024; }
// bytes of scratch the tail phase needs for (H, nsplit); the host checks it against the activation ring.
// The two H x 64 weight matrices do not live here: they arrive through the weight ring.
__host__ __device__ inline int tail_scratch_bytes(int H, int nsplit) {
  int need = H * (10 + 4 * nsplit);                                   // one final-layer row staged
  if (need < tail_n0_bytes(H)) need = tail_n0_bytes(H);               // next layer-0 tables
  return CH_TAIL_FIXED(H) + need;
}
// Consumer-side view of the weight ring (the MMA warp in GEMM phases, the epilogue warps in the step-boundary
// phase).  Every role walks the same program, so each can keep its own (slot, parity) cursor in step.
struct WRing {
  uint64_t* full; uint64_t* empty; uint8_t* base;
  int SB; int s; uint32_t p;
};
__device__ __forceinline__ void wring_advance(WRing& w, int k) {
  const int t = w.s + k;
  w.p ^= static_cast<uint32_t>(t / w.SB) & 1u;
  w.s = t % w.SB;
}
// weight-ring slots the step-boundary phase consumes in CTA `bid`: per owned row pair, w_out (H/64 tiles) and,
// unless it is the last step, w_in (H/64 tiles), CH_KPS tiles per slot
// (split: a pair is shared by CTAs 2q and 2q + 1, half of the tiles each; see ChainTail::split)
__device__ __forceinline__ int tail_w_slots(int H, int n, int last, int split, int bid, int grid) {
  if (split) {
    const int gh = grid >> 1, q = bid >> 1;
    const int pairs = (bid < 2 * gh && q < n) ? (n - q + gh - 1) / gh : 0;
    const int per = (H / 128 + CH_KPS - 1) / CH_KPS;
    return pairs * per * (last ? 1 : 2);
  }
  const int pairs = (bid < n) ? (n - bid + grid - 1) / grid : 0;
  const int per = (H / 64 + CH_KPS - 1) / CH_KPS;
  return pairs * per * (last ? 1 : 2);
}
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// A fragment (rows mt*16 + {g, g+8}, k-step ks) of a [64 x 64] bf16 K-major SWIZZLE_128B tile
__device__ __forceinline__ void ld_a_frag_sw128(const uint8_t* tile, int mt, int ks, int g, int t, uint32_t (&a)[4]) {
  const int row = mt * 16 + g;
  const uint8_t* r0 = tile + row * 128 + 4 * t;
  const uint8_t* r1 = r0 + 8 * 128;
  const int c0 = ((2 * ks) ^ (row & 7)) << 4, c1 = ((2 * ks + 1) ^ (row & 7)) << 4;   // (row + 8) & 7 == row & 7
  a[0] = *reinterpret_cast<const uint32_t*>(r0 + c0);
  a[1] = *reinterpret_cast<const uint32_t*>(r1 + c0);
  a[2] = *reinterpret_cast<const uint32_t*>(r0 + c1);
  a[3] = *reinterpret_cast<const uint32_t*>(r1 + c1);
}
__device__ __forceinline__ void phase_tail(const ChainTail& a, const ChainCall& call, RowCtx& c, WRing& wr, int ph) {
  constexpr int L = 64;
  const int H = a.H, n = a.n, nvec = H / 4, tid = c.tid, warp = tid >> 5, lane = tid & 31, M = 2 * n;
  const int g = lane >> 2, t = lane & 3;
  bf16* ybf = reinterpret_cast<bf16*>(c.scr);
  float* xs = reinterpret_cast<float*>(c.scr + 4 * H);
  float* vs = reinterpret_cast<float*>(c.scr + 8 * H);
  float* zs = vs + 2 * L;
  float* vsw = reinterpret_cast<float*>(c.scr + 8 * H + 1024);        // per-warp partial sums of the final Linear
  uint8_t* stg = c.scr + CH_TAIL_FIXED(H);                            // 1024-byte aligned (H % 128 == 0)
  const int stg_bytes = c.scr_bytes - CH_TAIL_FIXED(H);
  const int ntiles = H / 64, nslots = (ntiles + CH_KPS - 1) / CH_KPS;
  int nw = NUM_EPI_WARPS;                         // warps that consume weight-ring slots (see the final Linear)
  while (wr.SB % nw != 0) --nw;
#pragma unroll 1
  for (int p = blockIdx.x; p < n; p += gridDim.x) {
    // ---- final layer input of both rows: fold the last down-projection, RMSNorm, modulate ----
    const int rsb = (row_stage_bytes(H, a.nsplit) + 127) / 128 * 128;
    const bool both = 2 * rsb <= stg_bytes;           // both rows in flight together when the staging fits
    row_issue(c, 0, stg, a.x, a.partial, a.nsplit, M, p, a.prev_gate, a.fshift, a.fshift + H, a.ld_mod, H, a.rowmajor);
    if (both) row_issue(c, 1, stg + rsb, a.x, a.partial, a.nsplit, M, p + n, a.prev_gate, a.fshift, a.fshift + H, a.ld_mod, H, a.rowmajor);
#pragma unroll 1
    for (int r = 0; r < 2; ++r) {
      const int m = p + r * n;
      uint8_t* rs = stg + ((both && r == 1) ? rsb : 0);
      if (!both && r == 1) row_issue(c, 0, stg, a.x, a.partial, a.nsplit, M, m, a.prev_gate, a.fshift, a.fshift + H, a.ld_mod, H, a.rowmajor);
      const float ss = row_fold(c, (both && r == 1) ? 1 : 0, rs, a.nsplit, m, H, nullptr);
      const float rms = rsqrtf(block_sum256(c, ss, r) / static_cast<float>(H) + a.eps);
      const float* sx = reinterpret_cast<const float*>(rs);
      const bf16* ssh = reinterpret_cast<const bf16*>(rs + 6 * H);
      const bf16* ssc = ssh + H;
#pragma unroll 1
      for (int i = tid; i < nvec; i += CH_EPI_THREADS)
        *reinterpret_cast<uint2*>(ybf + r * H + i * 4) =
            modulate4(reinterpret_cast<const float4*>(sx)[i], rms, make_float4(1.f, 1.f, 1.f, 1.f),
                      reinterpret_cast<const uint2*>(ssh)[i], reinterpret_cast<const uint2*>(ssc)[i]);
      row_release(c);
    }
    if (tid == 0) CTRACE(16 * ph + 9, clock64());
    // ---- v[r][l] = yF[r] . w_out[l] : tile j = w_out[0:64, 64j : 64j+64]; warp w: rows 16*(w&3).., k-steps 2*(w>>2).. ----
    {
      // the tiles come through the weight ring: its producer warp loaded them while this CTA was still in the
      // previous phase / barrier, so the first ones are resident when the rows are ready
      // One warp per ring slot (slot q -> warp q % 8): a warp contracts whole tiles on its own and frees the slot
      // itself, so there is no block-wide barrier per slot (that loop ran at ~850 clk per slot whatever was
      // resident: wait + dependent mma.sync chain + barrier).  Partial sums meet in vs[] in a fixed order (below).
      // A physical slot must belong to ONE warp: a parity wait can only tell "phase k done" from "phase k+1 done"
      // if the waiter has seen phase k complete, so warp b must never wait for a slot's next fill while warp a is
      // still waiting for the current one.  nw = the largest divisor of the ring depth that is <= 8.
      float acc[4][4];
#pragma unroll
      for (int mt = 0; mt < 4; ++mt) { acc[mt][0] = 0.f; acc[mt][1] = 0.f; acc[mt][2] = 0.f; acc[mt][3] = 0.f; }
#pragma unroll 1
      for (int q = warp; q < nslots && warp < nw; q += nw) {
        const int idx = wr.s + q, slot = idx % wr.SB;
        mbar_wait(&wr.full[slot], wr.p ^ (static_cast<uint32_t>(idx / wr.SB) & 1u));
        const int cnt = min(CH_KPS, ntiles - q * CH_KPS);
#pragma unroll 1
        for (int j = 0; j < cnt; ++j) {
          const uint8_t* tile = wr.base + slot * CH_B_BYTES + j * CH_B_TILE;
          const int kbase = (q * CH_KPS + j) * 64;
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            uint32_t b0 = 0u, b1 = 0u;
            if (g < 2) {
              const bf16* yb = ybf + g * H + kbase + ks * 16 + 2 * t;
              b0 = *reinterpret_cast<const uint32_t*>(yb);
              b1 = *reinterpret_cast<const uint32_t*>(yb + 8);
            }
#pragma unroll
            for (int mt = 0; mt < 4; ++mt) {
              uint32_t af[4];
              ld_a_frag_sw128(tile, mt, ks, g, t, af);
              mma_bf16_16816(acc[mt], af, b0, b1);
            }
          }
        }
        __syncwarp();                                 // reads only: no proxy fence before the producer refills
        if (lane == 0) mbar_arrive(&wr.empty[slot]);
      }
      wring_advance(wr, nslots);
      // Partial sums meet in a FIXED order (warp 0 .. 7), so the sampled latent is bitwise reproducible: every warp
      // (idle ones with zeros) parks its 2 x 64 partials, then thread l < 128 adds the eight of them.
      if (t == 0) {                                   // columns 0/1 of the accumulators = the two rows
        float* mine = vsw + warp * (2 * L);
#pragma unroll
        for (int mt = 0; mt < 4; ++mt) {
          const int l = mt * 16 + g;
          mine[l] = acc[mt][0];
          mine[L + l] = acc[mt][1];
          mine[l + 8] = acc[mt][2];
          mine[L + l + 8] = acc[mt][3];
        }
      }
      epi_bar_sync();
      if (tid < 2 * L) {
        float sum = 0.f;
#pragma unroll
        for (int w = 0; w < NUM_EPI_WARPS; ++w) sum += vsw[w * (2 * L) + tid];
        vs[tid] = sum;
      }
      epi_bar_sync();
    }
    if (tid == 0) CTRACE(16 * ph + 10, clock64());
    // ---- CFG combine + v -> x0 + DPM-Solver++ update (k_cfg_update) ----
    if (tid < L) {
      const int i = p * L + tid;
      const int nL = n * L;
      const float cc = vs[tid], u = vs[L + tid];
      const float veps = u + call.cfg * (cc - u);
      const float x = __ldcg(a.zstate + i);
      const float x0 = a.alpha_s * x - a.sigma_s * veps;
      float xn = a.cx * x + a.c0 * x0;
      if (a.order == 2) xn += a.c1 * (a.inv_r0 * (x0 - __ldcg(a.prev_x0 + i)));
      if (a.cn != 0.f && call.sde_noise != nullptr) xn += a.cn * call.sde_noise[static_cast<size_t>(a.step) * nL + i];
      a.prev_x0[i] = x0;
      a.zstate[i] = xn;
      const bf16 bz = __float2bfloat16(xn);
      a.zb[i] = bz;
      a.zb[nL + i] = bz;
      reinterpret_cast<bf16*>(zs)[tid] = bz;
      if (a.last && call.out != nullptr) {
        if (call.out_dtype == VV_DTYPE_BF16) reinterpret_cast<bf16*>(call.out)[i] = bz;
        else reinterpret_cast<float*>(call.out)[i] = xn;
      }
    }
    epi_bar_sync();
    if (a.last) continue;
    if (tid == 0) CTRACE(16 * ph + 11, clock64());
    // ---- x = bf16(z') . w_in^T : tile j = w_in[64j : 64j+64, 0:64]; one (tile, 16-row slab) per warp pass ----
    uint8_t* n0 = stg;                                // next step's layer-0 tables: w0 [H f32] | row 0: shift0, scale0 [H bf16 each] | row 1: the same
    {
      if (tid == 0) mbar_arrive_expect_tx(&c.bars[0], 12 * H);
      if (lane == 0) {                                // three copies, one per warp (see row_issue); scale0 = shift0 + H
        if (warp == 0) bulk_g2s(n0, a.norm_w0, 4 * H, &c.bars[0]);
        if (warp == 1 || warp == 2) {
          const int r = warp - 1;
          bulk_g2s(n0 + 4 * H + r * 4 * H, a.shift0 + static_cast<size_t>(p + r * n) * a.ld_mod, 4 * H, &c.bars[0]);
        }
      }
      uint32_t zb0[4], zb1[4];                        // z as column 0 of the 8-wide operand, all four k-steps
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const bf16* zz = reinterpret_cast<const bf16*>(zs) + ks * 16 + 2 * t;
        zb0[ks] = (g == 0) ? *reinterpret_cast<const uint32_t*>(zz) : 0u;
        zb1[ks] = (g == 0) ? *reinterpret_cast<const uint32_t*>(zz + 8) : 0u;
      }
#pragma unroll 1
      for (int q = warp; q < nslots && warp < nw; q += nw) {   // w_in tiles, again one warp per weight-ring slot
        const int idx = wr.s + q, slot = idx % wr.SB;
        mbar_wait(&wr.full[slot], wr.p ^ (static_cast<uint32_t>(idx / wr.SB) & 1u));
        const int cnt = min(CH_KPS, ntiles - q * CH_KPS);
#pragma unroll 1
        for (int j = 0; j < cnt; ++j) {
          const uint8_t* tile = wr.base + slot * CH_B_BYTES + j * CH_B_TILE;
#pragma unroll
          for (int mt = 0; mt < 4; ++mt) {
            float d[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
              uint32_t af[4];
              ld_a_frag_sw128(tile, mt, ks, g, t, af);
              mma_bf16_16816(d, af, zb0[ks], zb1[ks]);
            }
            if (t == 0) {
              const int k = (q * CH_KPS + j) * 64 + mt * 16 + g;
              xs[k] = d[0];
              xs[k + 8] = d[2];
            }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&wr.empty[slot]);
      }
      wring_advance(wr, nslots);
      epi_bar_sync();
    }
    if (tid == 0) CTRACE(16 * ph + 12, clock64());
    // ---- layer-0 RMSNorm + modulate of the next step, both rows (same x, their own shift/scale) ----
    {
      float s2 = 0.f;
#pragma unroll 1
      for (int i = tid; i < nvec; i += CH_EPI_THREADS) {
        const float4 x4 = reinterpret_cast<const float4*>(xs)[i];
        s2 += x4.x * x4.x + x4.y * x4.y + x4.z * x4.z + x4.w * x4.w;
        reinterpret_cast<float4*>(a.x + static_cast<size_t>(p) * H)[i] = x4;
        reinterpret_cast<float4*>(a.x + static_cast<size_t>(p + n) * H)[i] = x4;
      }
      const float rms = rsqrtf(block_sum256(c, s2, 0) / static_cast<float>(H) + a.eps);
      row_wait(c, 0);
#pragma unroll 1
      for (int i = tid; i < nvec; i += CH_EPI_THREADS) {
        const float4 x4 = reinterpret_cast<const float4*>(xs)[i];
        const float4 wv = reinterpret_cast<const float4*>(n0)[i];
#pragma unroll
        for (int r = 0; r < 2; ++r)
          *reinterpret_cast<uint2*>(a.y + static_cast<size_t>(p + r * n) * H + i * 4) =
              modulate4(x4, rms, wv, reinterpret_cast<const uint2*>(n0 + 4 * H + r * 4 * H)[i],
                        reinterpret_cast<const uint2*>(n0 + 6 * H + r * 4 * H)[i]);
      }
    }
    row_release(c);
  }
}

// ---- the same step boundary with every row pair shared by two CTAs (ChainTail::split) ----
// phase_tail keeps 64 of 148 SMs busy for ~33 k clk, most of it waiting for the two H x 64 matrices to come through a
// 6-slot ring (196 KB each, per CTA) and for two 52 KB row sets to be staged.  Here CTA 2q + h owns columns
// [h * H/2, (h + 1) * H/2) of pair q: half of both folded rows, the final Linear over half of its K range, half of the
// input projection's outputs.  Each half of a matrix is 6 ring slots - resident before the rows are ready.  The halves
// meet three times through a mailbox in global memory (one 1 KB block per pair and half: 128 + 3 floats of data, three
// flag words), each time "write data, CTA barrier, one thread: fence + flag = tag" / "one thread: spin on the partner's
// flag, CTA barrier, read": (1) the two rows' sums of squares, (2) the final Linear's 2 x 64 partial sums, (3) the next
// norm's sum of squares.  Sums are always formed as (half 0's part) + (half 1's part): both CTAs get identical bits,
// and the result does not depend on anything but the pair's own data.  The tag is the phase's barrier epoch, unique
// over the program's life; a mailbox is reused one solver step (a dozen grid barriers) later at the earliest.
// Both CTAs run the CFG combine and the solver update (64 values); only half 0 stores the state.  Each reads the old
// state BEFORE it posts exchange (2), which half 0 must have received before it overwrites the state.
constexpr int CH_MAIL_FLOATS = 256;        // per (pair, half): data [0..127] | [128],[129] | [130] ; flags at words 192..194
__device__ __forceinline__ float* mail_of(float* mail, int p, int half) { return mail + (static_cast<size_t>(p) * 2 + half) * CH_MAIL_FLOATS; }
__device__ __forceinline__ void pair_post(float* mine, int k, uint32_t tag) {       // one thread, behind a CTA barrier
  __threadfence();
  st_release_u32(reinterpret_cast<uint32_t*>(mine) + 192 + k, tag);
}
__device__ __forceinline__ void pair_wait(const float* theirs, int k, uint32_t tag) {
  const long long t0 = clock64();
  while (ld_acquire_u32(reinterpret_cast<const uint32_t*>(theirs) + 192 + k) != tag) {
    if (clock64() - t0 > (1ll << 32)) {
      printf("vvb200: step-boundary exchange %d timed out (block %d, tag %u)\n", k, blockIdx.x, tag);
      __trap();
    }
  }
}
// column range [co, co + hl) of row m: x [hl f32] | gate [hl bf16] | shift [hl bf16] | scale [hl bf16] | partial [nsplit][hl f32]
__device__ __forceinline__ void rowc_issue(RowCtx& c, int bar, uint8_t* stg, const float* x, const float* partial, int nsplit,
                                           int M, int m, const bf16* gate, const bf16* shift, const bf16* scale, int ld_mod,
                                           int H, int co, int hl, int rowmajor) {
  if ((c.tid & 31) != 0) return;                             // one copy per warp, see row_issue
  const int w = c.tid >> 5;
  float* sx = reinterpret_cast<float*>(stg);
  bf16* sg = reinterpret_cast<bf16*>(stg + 4 * hl);
  float* sp = reinterpret_cast<float*>(stg + 10 * hl);
  uint64_t* b = &c.bars[bar];
  if (c.tid == 0) mbar_arrive_expect_tx(b, static_cast<uint32_t>(hl * (8 + (nsplit > 0 ? 2 + 4 * nsplit : 0))));
  for (int z = w; z < nsplit; z += NUM_EPI_WARPS)
    bulk_g2s(sp + static_cast<size_t>(z) * hl,
             partial + (rowmajor ? (static_cast<size_t>(m) * nsplit + z) * H : (static_cast<size_t>(z) * M + m) * H) + co, hl * 4, b);
  const int k0 = (nsplit > 0 ? nsplit : 0);
  if (w == (k0 + 0) % NUM_EPI_WARPS) bulk_g2s(sx, x + static_cast<size_t>(m) * H + co, hl * 4, b);
  if (w == (k0 + 1) % NUM_EPI_WARPS) bulk_g2s(sg + hl, shift + static_cast<size_t>(m) * ld_mod + co, hl * 2, b);
  if (w == (k0 + 2) % NUM_EPI_WARPS) bulk_g2s(sg + 2 * hl, scale + static_cast<size_t>(m) * ld_mod + co, hl * 2, b);
  if (nsplit > 0 && w == (k0 + 3) % NUM_EPI_WARPS) bulk_g2s(sg, gate + static_cast<size_t>(m) * ld_mod + co, hl * 2, b);
}
__device__ __forceinline__ void phase_tail_split(const ChainTail& a, const ChainCall& call, RowCtx& c, WRing& wr, int ph, uint32_t tag) {
  constexpr int L = 64;
  const int H = a.H, n = a.n, hl = H >> 1, nvec = hl / 4, tid = c.tid, warp = tid >> 5, lane = tid & 31, M = 2 * n;
  const int g = lane >> 2, t = lane & 3;
  const int gh = gridDim.x >> 1, half = blockIdx.x & 1, co = half * hl;
  if (static_cast<int>(blockIdx.x) >= 2 * gh) return;
  bf16* ybf = reinterpret_cast<bf16*>(c.scr);                         // [2][hl]
  float* xs = reinterpret_cast<float*>(c.scr + 4 * H);                // [hl]
  float* vs = reinterpret_cast<float*>(c.scr + 8 * H);                // [2][64]
  float* zs = vs + 2 * L;
  float* vsw = reinterpret_cast<float*>(c.scr + 8 * H + 1024);        // per-warp partial sums of the final Linear
  uint8_t* stg = c.scr + CH_TAIL_FIXED(H);
  const int ntiles = hl / 64, nslots = (ntiles + CH_KPS - 1) / CH_KPS;
  int nw = NUM_EPI_WARPS;
  while (wr.SB % nw != 0) --nw;
  const int rsb = (row_stage_bytes(hl, a.nsplit) + 127) / 128 * 128;
#pragma unroll 1
  for (int p = blockIdx.x >> 1; p < n; p += gh) {
    float* mine = mail_of(a.mail, p, half);
    const float* theirs = mail_of(a.mail, p, half ^ 1);
    // the solver's old state (see the header comment: before anything is posted)
    float zx = 0.f, zp = 0.f;
    if (tid < L) {
      zx = __ldcg(a.zstate + p * L + tid);
      if (a.order == 2) zp = __ldcg(a.prev_x0 + p * L + tid);
    }
    if (tid == 0) CTRACE2(ph, 1, 0, clock64());
    // ---- final layer input, both rows, own columns: fold the last down-projection ----
    rowc_issue(c, 0, stg, a.x, a.partial, a.nsplit, M, p, a.prev_gate, a.fshift, a.fshift + H, a.ld_mod, H, co, hl, a.rowmajor);
    rowc_issue(c, 1, stg + rsb, a.x, a.partial, a.nsplit, M, p + n, a.prev_gate, a.fshift, a.fshift + H, a.ld_mod, H, co, hl, a.rowmajor);
    if (tid == 0) CTRACE2(ph, 1, 1, clock64());
    float ssr[2];
#pragma unroll 1
    for (int r = 0; r < 2; ++r) ssr[r] = block_sum256(c, row_fold(c, r, stg + r * rsb, a.nsplit, p + r * n, hl, nullptr), r);
    if (tid == 0) CTRACE2(ph, 1, 3, clock64());
    if (tid == 0) {                                                    // exchange (1)
      mine[128] = ssr[0];
      mine[129] = ssr[1];
      pair_post(mine, 0, tag);
      CTRACE2(ph, 1, 4, clock64());
      pair_wait(theirs, 0, tag);
      CTRACE2(ph, 1, 5, clock64());
      vs[0] = __ldcg(theirs + 128);
      vs[1] = __ldcg(theirs + 129);
    }
    epi_bar_sync();
#pragma unroll 1
    for (int r = 0; r < 2; ++r) {
      const float tot = half == 0 ? ssr[r] + vs[r] : vs[r] + ssr[r];
      const float rms = rsqrtf(tot / static_cast<float>(H) + a.eps);
      const uint8_t* rs = stg + r * rsb;
      const float* sx = reinterpret_cast<const float*>(rs);
      const bf16* ssh = reinterpret_cast<const bf16*>(rs + 6 * hl);
      const bf16* ssc = ssh + hl;
#pragma unroll 1
      for (int i = tid; i < nvec; i += CH_EPI_THREADS)
        *reinterpret_cast<uint2*>(ybf + r * hl + i * 4) =
            modulate4(reinterpret_cast<const float4*>(sx)[i], rms, make_float4(1.f, 1.f, 1.f, 1.f),
                      reinterpret_cast<const uint2*>(ssh)[i], reinterpret_cast<const uint2*>(ssc)[i]);
    }
    row_release(c);
    if (tid == 0) CTRACE(16 * ph + 9, clock64());
    if (tid == 0) CTRACE2(ph, 1, 6, clock64());
    // ---- v[r][l] partial over the own K range: tile j = w_out[0:64, co + 64j : co + 64j + 64] (phase_tail's scheme) ----
    float vsum = 0.f;
    {
      float acc[4][4];
#pragma unroll
      for (int mt = 0; mt < 4; ++mt) { acc[mt][0] = 0.f; acc[mt][1] = 0.f; acc[mt][2] = 0.f; acc[mt][3] = 0.f; }
#pragma unroll 1
      for (int q = warp; q < nslots && warp < nw; q += nw) {
        const int idx = wr.s + q, slot = idx % wr.SB;
        mbar_wait(&wr.full[slot], wr.p ^ (static_cast<uint32_t>(idx / wr.SB) & 1u));
        if (tid == 0) CTRACE2(ph, 1, 7, clock64());
        const int cnt = min(CH_KPS, ntiles - q * CH_KPS);
#pragma unroll 1
        for (int j = 0; j < cnt; ++j) {
          const uint8_t* tile = wr.base + slot * CH_B_BYTES + j * CH_B_TILE;
          const int kbase = (q * CH_KPS + j) * 64;
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            uint32_t b0 = 0u, b1 = 0u;
            if (g < 2) {
              const bf16* yb = ybf + g * hl + kbase + ks * 16 + 2 * t;
              b0 = *reinterpret_cast<const uint32_t*>(yb);
              b1 = *reinterpret_cast<const uint32_t*>(yb + 8);
            }
#pragma unroll
            for (int mt = 0; mt < 4; ++mt) {
              uint32_t af[4];
              ld_a_frag_sw128(tile, mt, ks, g, t, af);
              mma_bf16_16816(acc[mt], af, b0, b1);
            }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&wr.empty[slot]);
      }
      if (tid == 0) CTRACE2(ph, 1, 8, clock64());
      wring_advance(wr, nslots);
      if (t == 0) {
        float* w_ = vsw + warp * (2 * L);
#pragma unroll
        for (int mt = 0; mt < 4; ++mt) {
          const int l = mt * 16 + g;
          w_[l] = acc[mt][0];
          w_[L + l] = acc[mt][1];
          w_[l + 8] = acc[mt][2];
          w_[L + l + 8] = acc[mt][3];
        }
      }
      epi_bar_sync();
      if (tid < 2 * L) {
#pragma unroll
        for (int w = 0; w < NUM_EPI_WARPS; ++w) vsum += vsw[w * (2 * L) + tid];
        mine[tid] = vsum;
      }
      epi_bar_sync();
      if (tid == 0) {                                                  // exchange (2)
        CTRACE2(ph, 1, 9, clock64());
        pair_post(mine, 1, tag);
        pair_wait(theirs, 1, tag);
        CTRACE2(ph, 1, 10, clock64());
      }
      epi_bar_sync();
      if (tid < 2 * L) {
        const float o = __ldcg(theirs + tid);
        vs[tid] = half == 0 ? vsum + o : o + vsum;
      }
      epi_bar_sync();
    }
    if (tid == 0) CTRACE(16 * ph + 10, clock64());
    // ---- CFG combine + v -> x0 + DPM-Solver++ update (both CTAs compute it; half 0 stores the state) ----
    if (tid < L) {
      const int i = p * L + tid;
      const int nL = n * L;
      const float cc = vs[tid], u = vs[L + tid];
      const float veps = u + call.cfg * (cc - u);
      const float x0 = a.alpha_s * zx - a.sigma_s * veps;
      float xn = a.cx * zx + a.c0 * x0;
      if (a.order == 2) xn += a.c1 * (a.inv_r0 * (x0 - zp));
      if (a.cn != 0.f && call.sde_noise != nullptr) xn += a.cn * call.sde_noise[static_cast<size_t>(a.step) * nL + i];
      const bf16 bz = __float2bfloat16(xn);
      reinterpret_cast<bf16*>(zs)[tid] = bz;
      if (half == 0) {
        a.prev_x0[i] = x0;
        a.zstate[i] = xn;
        a.zb[i] = bz;
        a.zb[nL + i] = bz;
        if (a.last && call.out != nullptr) {
          if (call.out_dtype == VV_DTYPE_BF16) reinterpret_cast<bf16*>(call.out)[i] = bz;
          else reinterpret_cast<float*>(call.out)[i] = xn;
        }
      }
    }
    epi_bar_sync();
    if (a.last) continue;
    if (tid == 0) CTRACE(16 * ph + 11, clock64());
    if (tid == 0) CTRACE2(ph, 1, 11, clock64());
    // ---- x[co + k] = bf16(z') . w_in[co + k, :] : tile j = w_in[co + 64j : co + 64j + 64, 0:64] ----
    uint8_t* n0 = stg;                                // next step's layer-0 tables, own columns: w0 [hl f32] | shift0 [2][hl] | scale0 [2][hl]
    {
      if (tid == 0) mbar_arrive_expect_tx(&c.bars[0], 12 * hl);
      if (lane == 0) {                                // five copies, one per warp (see row_issue)
        if (warp == 0) bulk_g2s(n0, a.norm_w0 + co, 4 * hl, &c.bars[0]);
        if (warp >= 1 && warp <= 4) {
          const int r = (warp - 1) & 1, sc = (warp - 1) >> 1;          // shift0 / scale0 of row r
          const size_t m = static_cast<size_t>(p + r * n);
          bulk_g2s(n0 + (sc ? 8 : 4) * hl + r * 2 * hl, a.shift0 + (sc ? H : 0) + m * a.ld_mod + co, 2 * hl, &c.bars[0]);
        }
      }
      uint32_t zb0[4], zb1[4];
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const bf16* zz = reinterpret_cast<const bf16*>(zs) + ks * 16 + 2 * t;
        zb0[ks] = (g == 0) ? *reinterpret_cast<const uint32_t*>(zz) : 0u;
        zb1[ks] = (g == 0) ? *reinterpret_cast<const uint32_t*>(zz + 8) : 0u;
      }
#pragma unroll 1
      for (int q = warp; q < nslots && warp < nw; q += nw) {
        const int idx = wr.s + q, slot = idx % wr.SB;
        mbar_wait(&wr.full[slot], wr.p ^ (static_cast<uint32_t>(idx / wr.SB) & 1u));
        const int cnt = min(CH_KPS, ntiles - q * CH_KPS);
#pragma unroll 1
        for (int j = 0; j < cnt; ++j) {
          const uint8_t* tile = wr.base + slot * CH_B_BYTES + j * CH_B_TILE;
#pragma unroll
          for (int mt = 0; mt < 4; ++mt) {
            float d[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
              uint32_t af[4];
              ld_a_frag_sw128(tile, mt, ks, g, t, af);
              mma_bf16_16816(d, af, zb0[ks], zb1[ks]);
            }
            if (t == 0) {
              const int k = (q * CH_KPS + j) * 64 + mt * 16 + g;
              xs[k] = d[0];
              xs[k + 8] = d[2];
            }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&wr.empty[slot]);
      }
      wring_advance(wr, nslots);
      epi_bar_sync();
    }
    if (tid == 0) CTRACE(16 * ph + 12, clock64());
    if (tid == 0) CTRACE2(ph, 1, 13, clock64());
    // ---- layer-0 RMSNorm + modulate of the next step, both rows, own columns ----
    {
      float s2 = 0.f;
#pragma unroll 1
      for (int i = tid; i < nvec; i += CH_EPI_THREADS) {
        const float4 x4 = reinterpret_cast<const float4*>(xs)[i];
        s2 += x4.x * x4.x + x4.y * x4.y + x4.z * x4.z + x4.w * x4.w;
        reinterpret_cast<float4*>(a.x + static_cast<size_t>(p) * H + co)[i] = x4;
        reinterpret_cast<float4*>(a.x + static_cast<size_t>(p + n) * H + co)[i] = x4;
      }
      const float s2h = block_sum256(c, s2, 0);
      if (tid == 0) {                                                  // exchange (3)
        mine[130] = s2h;
        CTRACE2(ph, 1, 12, clock64());
        pair_post(mine, 2, tag);
        pair_wait(theirs, 2, tag);
        CTRACE2(ph, 1, 14, clock64());
        vs[0] = __ldcg(theirs + 130);
      }
      epi_bar_sync();
      const float tot = half == 0 ? s2h + vs[0] : vs[0] + s2h;
      const float rms = rsqrtf(tot / static_cast<float>(H) + a.eps);
      row_wait(c, 0);
#pragma unroll 1
      for (int i = tid; i < nvec; i += CH_EPI_THREADS) {
        const float4 x4 = reinterpret_cast<const float4*>(xs)[i];
        const float4 wv = reinterpret_cast<const float4*>(n0)[i];
#pragma unroll
        for (int r = 0; r < 2; ++r)
          *reinterpret_cast<uint2*>(a.y + static_cast<size_t>(p + r * n) * H + co + i * 4) =
              modulate4(x4, rms, wv, reinterpret_cast<const uint2*>(n0 + 4 * hl + r * 2 * hl)[i],
                        reinterpret_cast<const uint2*>(n0 + 8 * hl + r * 2 * hl)[i]);
      }
    }
    row_release(c);
    if (tid == 0) CTRACE2(ph, 1, 15, clock64());
  }
}

// work items of GEMM phase P owned by this CTA: item = blockIdx.x + i * gridDim.x
__device__ __forceinline__ int kb_range(const ChainPhase& P, int item, int* kb0) {
  const int z = item / (P.n_tiles_n * P.n_tiles_m);
  const int kb_total = (P.g.K + BLOCK_K - 1) / BLOCK_K;
  *kb0 = z * P.g.kb_per_split;
  const int kb1 = min(kb_total, *kb0 + P.g.kb_per_split);
  return max(kb1 - *kb0, 0);
}
__device__ __forceinline__ int kb_range(const ChainLite& L, int item, int* kb0) {
  const int z = item / (static_cast<int>(L.n_tiles_n) * L.n_tiles_m);
  *kb0 = z * L.kb_per_split;
  const int kb1 = min(static_cast<int>(L.kb_total), *kb0 + L.kb_per_split);
  return max(kb1 - *kb0, 0);
}

// Every CTA of a GEMM phase reads the same activation tiles.  Walking K in the same order would make all
// 148 CTAs ask L2 for the same 128 lines at the same moment; each work item therefore starts at its own
// k-block and wraps around (fp32 accumulation order differs per tile, deterministically).
__device__ __forceinline__ int k_rotation(int krot, int n_blk, int nkb) {
  return (krot != 0 && nkb > 1) ? (n_blk * krot) % nkb : 0;
}

// Two instantiations: STAGE = false walks the head's programs (GEMM, norm, step-boundary and CFG phases), STAGE = true
// the decoder / encoder stage programs (GEMM and mixer phases).  One kernel for both kept every row phase's registers
// in one allocation: pressure in a mixer phase spilled values on the head's hot paths (head 0.90 -> 1.02 ms when the
// mixer's loads were batched).
template <bool STAGE>
__global__ void __launch_bounds__(CH_THREADS, 1)
k_chain(const ChainPhase* __restrict__ prog_g, const ChainLite* __restrict__ lite, int n_phases, int SA, int SB,
        const CUtensorMap* __restrict__ tmaps, const ChainCall call, uint32_t* flags, uint32_t epoch_base, int krot, int push) {
  extern __shared__ uint8_t smem_raw[];
  // align by offset (not by integer round trip) so the compiler keeps the shared address space
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + SA * CH_A_BYTES;
  ChainSmem* cs = reinterpret_cast<ChainSmem*>(smem_b + SB * CH_B_BYTES);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < SA; ++s) {
      mbar_init(&cs->full_a[s], 1);
      mbar_init(&cs->empty_a[s], 1);
    }
    for (int s = 0; s < SB; ++s) {
      mbar_init(&cs->full_b[s], 1);
      mbar_init(&cs->empty_b[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&cs->acc_full[b], 2);                // one arrival per MMA issuer
      mbar_init(&cs->acc_empty[b], NUM_EPI_WARPS);
      mbar_init(&cs->win_bar[b], 1);
    }
    for (int b = 0; b < CH_ROW_BARS; ++b) mbar_init(&cs->row_bar[b], 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(&cs->tmem_slot, CH_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = cs->tmem_slot;

  if (warp < CH_EPI_WARP0) {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(CH_REGS_LIGHT));
  if (warp == CH_W_WARP) {
    // ------------------------------ weight-tile producer ----------------------
    // Walks every GEMM phase of the program in order and keeps the weight ring full.  It never
    // waits for a grid barrier: while the other roles sit in a barrier or a row phase, the weight
    // tiles of the coming GEMM phases keep arriving, so HBM streams through the whole launch and a
    // GEMM phase only has to wait for its activation tiles (L2 hits) after its barrier.
    // (An L2 prefetch of the following phase's tiles from here - cp.async.bulk.prefetch.tensor - was measured twice,
    // with one and with two MMA issuers: 1.14-1.17 ms against 1.09, 1.06-1.10 against 1.04.  Removed.)
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      ChainLite nxt = lite_ld(lite, 0);
      for (int ph = 0; ph < n_phases; ++ph) {
        const ChainLite L = nxt;
        if (ph + 1 < n_phases) nxt = lite_ld(lite, ph + 1);
        if (!STAGE && L.op == CH_TAIL) {
          // the step-boundary phase's two H x 64 matrices, as [64 x 64] tiles: w_out tile j = (k-block j, N tile 0),
          // w_in tile j = (k-block 0, N tile j); consumed by the epilogue warps (phase_tail)
          const int tn = L.n_items, tlast = L.flag, split = L.n_tiles_m;
          const CUtensorMap* tmO = &tmaps[L.tmA];
          const CUtensorMap* tmI = &tmaps[L.tmB];
          tma_prefetch_desc(tmO);
          if (!tlast) tma_prefetch_desc(tmI);
          // split: CTAs 2q and 2q + 1 share pair q, the even one takes the first half of the tiles
          const int gh = gridDim.x >> 1;
          const int ntiles = split ? L.kb_total >> 1 : L.kb_total;
          const int tile0 = split ? (blockIdx.x & 1) * ntiles : 0;
          const int pr0 = split ? ((static_cast<int>(blockIdx.x) < 2 * gh) ? static_cast<int>(blockIdx.x >> 1) : tn) : static_cast<int>(blockIdx.x);
          const int prs = split ? gh : static_cast<int>(gridDim.x);
          for (int pr = pr0; pr < tn; pr += prs) {
            for (int mat = 0; mat < (tlast ? 1 : 2); ++mat) {
              for (int t0 = 0; t0 < ntiles; t0 += CH_KPS) {
                const int cnt = min(CH_KPS, ntiles - t0);
                mbar_wait(&cs->empty_b[stage], phase ^ 1);
                mbar_arrive_expect_tx(&cs->full_b[stage], static_cast<uint32_t>(cnt * CH_B_TILE));
                for (int i = 0; i < cnt; ++i) {
                  uint8_t* dst = smem_b + stage * CH_B_BYTES + i * CH_B_TILE;
                  if (mat == 0) tma_load_2d(dst, tmO, &cs->full_b[stage], (tile0 + t0 + i) * BLOCK_K, 0, VV_L2_EVICT_LAST);
                  else tma_load_2d(dst, tmI, &cs->full_b[stage], 0, (tile0 + t0 + i) * CH_BN, VV_L2_EVICT_LAST);
                }
                if (++stage == SB) { stage = 0; phase ^= 1; }
              }
            }
          }
          continue;
        }
        if (L.op != CH_GEMM) continue;
        const CUtensorMap* tmB = &tmaps[L.tmB];
        const unsigned long long pol = L.flag ? VV_L2_EVICT_LAST : VV_L2_EVICT_FIRST;
        tma_prefetch_desc(tmB);
        for (int item = blockIdx.x; item < L.n_items; item += gridDim.x) {
          const int n_blk = item % L.n_tiles_n;
          int kb0;
          const int nkb = kb_range(L, item, &kb0);
          const int ngr = (nkb + CH_KPS - 1) / CH_KPS;
          int gr = k_rotation(krot, n_blk, ngr);
          for (int j = 0; j < ngr; ++j, gr = (gr + 1 == ngr) ? 0 : gr + 1) {
            const int cnt = min(CH_KPS, nkb - gr * CH_KPS);
            mbar_wait(&cs->empty_b[stage], phase ^ 1);
            CTRACE2(ph, 2, j, clock64());
            mbar_arrive_expect_tx(&cs->full_b[stage], static_cast<uint32_t>(cnt * CH_B_TILE));
            for (int i = 0; i < cnt; ++i)
              tma_load_2d(smem_b + stage * CH_B_BYTES + i * CH_B_TILE, tmB, &cs->full_b[stage],
                          (kb0 + gr * CH_KPS + i) * BLOCK_K, n_blk * CH_BN, pol);
            if (++stage == SB) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 0) {
    // ------------------------------ activation-tile producer, grid-barrier leader ------------------
    // Everything the first activation group of a GEMM phase needs is done BEFORE the barrier that opens the phase
    // (descriptor prefetch, the work item's coordinates, the wait for the ring slot, arming its mbarrier), so that the
    // TMA goes out the moment the leader has seen the barrier complete.
    struct Plan { int kb0, nkb, ngr, m_row0, gr; };
    auto plan = [&](const ChainLite& L, int item) {
      Plan q;
      q.nkb = kb_range(L, item, &q.kb0);
      q.ngr = (q.nkb + CH_KPS - 1) / CH_KPS;
      q.m_row0 = ((item / L.n_tiles_n) % L.n_tiles_m) * BLOCK_M;
      q.gr = k_rotation(krot, item % L.n_tiles_n, q.ngr);
      return q;
    };
    int stage = 0;
    uint32_t phase = 0;
    ChainLite nxt = lite_ld(lite, 0);
    Plan first{};
    bool armed = false;                         // the coming phase's first group: slot waited for, barrier armed
    for (int ph = 0; ph < n_phases; ++ph) {
      const ChainLite L = nxt;
      if (ph + 1 < n_phases) nxt = lite_ld(lite, ph + 1);     // lands while this phase runs: no L2 trip behind the barrier
      if (lane == 0 && L.op == CH_GEMM) {
        const CUtensorMap* tmA = &tmaps[L.tmA];
        const uint32_t a_bytes = static_cast<uint32_t>(L.a_rows * BLOCK_K * 2);
        for (int item = blockIdx.x; item < L.n_items; item += gridDim.x) {
          const Plan q = (item == static_cast<int>(blockIdx.x) && ph > 0) ? first : plan(L, item);
          int gr = q.gr;
          for (int j = 0; j < q.ngr; ++j, gr = (gr + 1 == q.ngr) ? 0 : gr + 1) {
            const int cnt = min(CH_KPS, q.nkb - gr * CH_KPS);
            if (!armed) {
              mbar_wait(&cs->empty_a[stage], phase ^ 1);
              mbar_arrive_expect_tx(&cs->full_a[stage], a_bytes * cnt);
            }
            armed = false;
            for (int i = 0; i < cnt; ++i)
              tma_load_2d(smem_a + stage * CH_A_BYTES + i * CH_A_TILE, tmA, &cs->full_a[stage],
                          (q.kb0 + gr * CH_KPS + i) * BLOCK_K, q.m_row0, VV_L2_EVICT_NORMAL);
            if (++stage == SA) { stage = 0; phase ^= 1; }
          }
        }
      }
      if (lane == 0 && ph + 1 < n_phases && nxt.op == CH_GEMM) {
        tma_prefetch_desc(&tmaps[nxt.tmA]);
        if (static_cast<int>(blockIdx.x) < nxt.n_items) {
          first = plan(nxt, blockIdx.x);
          if (first.ngr > 0) {
            // the slot is free once this phase's MMAs have completed, which the barrier below waits for anyway
            mbar_wait(&cs->empty_a[stage], phase ^ 1);
            mbar_arrive_expect_tx(&cs->full_a[stage],
                                  static_cast<uint32_t>(nxt.a_rows * BLOCK_K * 2) * min(CH_KPS, first.nkb - first.gr * CH_KPS));
            armed = true;
          }
        }
      }
      __syncwarp();
      grid_sync<true>(flags, epoch_base + static_cast<uint32_t>(ph) + 1u, ph, push);
    }
  } else if (warp == 1 || warp == CH_MMA2_WARP) {
    // ------------------------------ MMA issuers --------------------------------
    // The whole warp walks the loop converged; elect.sync picks the issuing lane (umma_*_warp, vv_common.cuh).
    // Both warps walk every ring group of every work item and keep the same ring cursors; warp `w` issues the groups
    // whose activation-ring slot has parity w - the running group index, since the ring depths are even
    // (chain_finalize) - into accumulator half w of the item's TMEM buffer.  Within a phase a ring slot is thus always
    // waited on by the same warp.  That matters: a parity wait cannot tell "this fill" from "two fills ago", and a
    // second waiter that ran ahead of the first by a whole ring would take a stale slot for a fresh one (seen as
    // sporadic wrong results when ownership went by the group's index within its work item).  The weight ring's
    // cursor can shift against the activation ring's by an odd count in a step-boundary phase; ownership changes
    // then happen across a CTA barrier, where both warps have consumed everything before it.
    constexpr uint32_t idesc = umma_idesc_bf16(BLOCK_M, CH_BN);
    const int w = __shfl_sync(0xffffffffu, warp == 1 ? 0 : 1, 0);     // provably warp-uniform (see below)
    int sa = 0, sb = 0;
    uint32_t pa = 0, pb = 0;
    uint32_t acc_cnt = 0;
    ChainLite nxt = lite_ld(lite, 0);
    for (int ph = 0; ph < n_phases; ++ph) {
      // loop bounds must be provably warp-uniform (loaded values are not): otherwise ptxas treats the loop body
      // as divergent and wraps every UTCHMMA in an ELECT / BRA.U.ANY loop
      const ChainLite L = lite_uniform(nxt);
      if (ph + 1 < n_phases) nxt = lite_ld(lite, ph + 1);
      const int op = L.op;
      if (op == CH_GEMM) {
        const int n_items = L.n_items;
        const int n_tiles_n = L.n_tiles_n;
        const int n_tiles_nm = n_tiles_n * L.n_tiles_m;
        const int kb_total = L.kb_total;
        const int kb_per_split = L.kb_per_split;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++acc_cnt) {
          const int kb0 = (item / n_tiles_nm) * kb_per_split;
          const int nkb = max(min(kb_total, kb0 + kb_per_split) - kb0, 0);
          const uint32_t buf = acc_cnt & 1;
          mbar_wait(&cs->acc_empty[buf], ((acc_cnt >> 1) & 1) ^ 1);   // epilogue drained this buffer
          tc_fence_after();
          const uint32_t d = tmem_base + buf * (2 * CH_BN) + w * CH_BN;
#ifdef VV_GEMM_TRACE
          long long wait_a = 0, wait_b = 0;
#endif
          const int ngr = (nkb + CH_KPS - 1) / CH_KPS;
          int gr = k_rotation(krot, item % n_tiles_n, ngr);
          int mine = 0;                                              // groups this warp has issued for the item
          for (int j = 0; j < ngr; ++j, gr = (gr + 1 == ngr) ? 0 : gr + 1) {
            if ((sa & 1) == w) {
              const int cnt = min(CH_KPS, nkb - gr * CH_KPS);
#ifdef VV_GEMM_TRACE
              const long long tw0 = clock64();
#endif
              mbar_wait(&cs->full_b[sb], pb);
#ifdef VV_GEMM_TRACE
              const long long tw1 = clock64();
#endif
              mbar_wait(&cs->full_a[sa], pa);
              tc_fence_after();
#ifdef VV_GEMM_TRACE
              const long long tw2 = clock64();
              wait_b += tw1 - tw0; wait_a += tw2 - tw1;
              if (lane == 0) {
                CTRACE2(ph, 0, j, tw1); CTRACE2(ph, 1, j, tw2);
                if (j == 0) { CTRACE(16 * ph + 13, tw1); CTRACE(16 * ph + 8, tw2); }
                if (j == ngr - 1) { CTRACE(16 * ph + 11, tw2); CTRACE(16 * ph + 9, wait_b); CTRACE(16 * ph + 10, wait_a); }
              }
#endif
              const uint64_t adesc = umma_desc_sw128(smem_u32(smem_a + sa * CH_A_BYTES));
              const uint64_t bdesc = umma_desc_sw128(smem_u32(smem_b + sb * CH_B_BYTES));
#pragma unroll
              for (int i = 0; i < CH_KPS; ++i) {
                if (i < cnt) {
#pragma unroll
                  for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
                    umma_bf16_warp(d, adesc + i * (CH_A_TILE >> 4) + 2 * k, bdesc + i * (CH_B_TILE >> 4) + 2 * k, idesc,
                                   (mine | i | k) != 0);
                }
              }
              umma_commit_warp(&cs->empty_a[sa]);
              umma_commit_warp(&cs->empty_b[sb]);
              ++mine;
            }
            if (++sa == SA) { sa = 0; pa ^= 1; }
            if (++sb == SB) { sb = 0; pb ^= 1; }
          }
          // one arrival per issuer: behind its MMAs, or a plain one when it had no group (an item of one group)
          if (mine > 0) umma_commit_warp(&cs->acc_full[buf]);
          else if (lane == 0) mbar_arrive(&cs->acc_full[buf]);
          if (lane == 0 && w == 0) CTRACE(16 * ph + 12, clock64());
        }
      } else if (!STAGE && op == CH_TAIL) {
        // the epilogue warps consume weight-ring slots in this phase: keep this warp's cursor in step
        const int k = tail_w_slots(L.kb_total * 64, L.n_items, L.flag, L.n_tiles_m, blockIdx.x, gridDim.x);
        const int tt = sb + k;
        pb ^= static_cast<uint32_t>(tt / SB) & 1u;
        sb = tt % SB;
      }
      grid_sync<false>(flags, epoch_base + static_cast<uint32_t>(ph) + 1u, ph, push);
    }
  }
  } else {
    // ------------------------------ epilogue / row-phase warps -----------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(CH_REGS_HEAVY));
    const int tid = threadIdx.x - 32 * CH_EPI_WARP0;
    const int q = warp & 3;                           // TMEM lanes 32 q .. 32 q + 31 are the ones this warp may read
    const int half = (warp - CH_EPI_WARP0) >> 2;
    uint32_t acc_cnt = 0;
    uint32_t agrp = 0;                                // ring groups of the activation ring so far (its parity picks the issuer)
    RowCtx rc;
    rc.bars = cs->row_bar; rc.parity = 0u; rc.scr = smem_a; rc.scr_bytes = SA * CH_A_BYTES; rc.red = cs->red; rc.tid = tid;
    WRing wr;
    wr.full = cs->full_b; wr.empty = cs->empty_b; wr.base = smem_b; wr.SB = SB; wr.s = 0; wr.p = 0u;
    // full program entries come through a two-slot window: entry ph + 1 is fetched (bulk copy) when phase ph starts
    if (tid == 0) {
      mbar_arrive_expect_tx(&cs->win_bar[0], static_cast<uint32_t>(sizeof(ChainPhase)));
      bulk_g2s(&cs->win[0], prog_g, static_cast<uint32_t>(sizeof(ChainPhase)), &cs->win_bar[0]);
    }
    for (int ph = 0; ph < n_phases; ++ph) {
      mbar_wait(&cs->win_bar[ph & 1], static_cast<uint32_t>(ph >> 1) & 1u);
      if (tid == 0 && ph + 1 < n_phases) {
        // the other slot held entry ph - 1: every epilogue thread was done with it before the last grid barrier
        fence_proxy_async_smem();
        mbar_arrive_expect_tx(&cs->win_bar[(ph + 1) & 1], static_cast<uint32_t>(sizeof(ChainPhase)));
        bulk_g2s(&cs->win[(ph + 1) & 1], prog_g + ph + 1, static_cast<uint32_t>(sizeof(ChainPhase)), &cs->win_bar[(ph + 1) & 1]);
      }
      const ChainPhase& P = cs->win[ph & 1];
      rc.ph = ph;
      if (P.op == CH_GEMM) {
        const GemmKParams g = P.g;
        for (int item = blockIdx.x; item < P.n_items; item += gridDim.x, ++acc_cnt) {
          const int n_blk = item % P.n_tiles_n, m_blk = (item / P.n_tiles_n) % P.n_tiles_m, z = item / (P.n_tiles_n * P.n_tiles_m);
          int kb0_;
          const int ngr_item = (kb_range(P, item, &kb0_) + CH_KPS - 1) / CH_KPS;
          wring_advance(wr, ngr_item);                           // slots the MMA warps consume for this item
          const uint32_t first_acc = (ngr_item == 1) ? (agrp & 1u) : 0u;   // a one-group item sits in its issuer's accumulator
          agrp += static_cast<uint32_t>(ngr_item);
          const uint32_t buf = acc_cnt & 1;
          mbar_wait(&cs->acc_full[buf], (acc_cnt >> 1) & 1);
          tc_fence_after();
          if (tid == 0) CTRACE(16 * ph + 14, clock64());
          uint32_t r[32];
          const uint32_t tacc = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * (2 * CH_BN) + half * 32;
          tmem_ld_32x32(tacc + first_acc * CH_BN, r);
          tmem_ld_wait();
          if (ngr_item > 1) {                         // two or more ring groups: both issuers' accumulators hold a part
#pragma unroll
            for (int h2 = 0; h2 < 2; ++h2) {
              uint32_t r2[16];
              tmem_ld_16(tacc + CH_BN + h2 * 16, r2);
              tmem_ld_wait();
#pragma unroll
              for (int k = 0; k < 16; ++k) r[h2 * 16 + k] = __float_as_uint(__uint_as_float(r[h2 * 16 + k]) + __uint_as_float(r2[k]));
            }
          }
          if (tid == 0) CTRACE(16 * ph + 15, clock64());
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&cs->acc_empty[buf]);   // values are in registers: release the buffer
          if (tid == 0) CTRACE2(ph, 0, 12, clock64());
          const int row = m_blk * BLOCK_M + q * 32 + lane;
          const int col0 = n_blk * CH_BN + half * 32;
          if (g.mode == GM_PARTIAL_F32 && P.n_items <= static_cast<int>(gridDim.x) && col0 + 32 <= g.N && (g.ldo & 3) == 0) {
            // Split-K partials: after tcgen05.ld a lane owns a ROW, so storing its 32 fp32 directly is eight
            // instructions of 32 half-filled sector writes each (3.9 k clk per phase, r01_mma_issue.md).  Transpose
            // the warp's 32 x 32 block through shared memory and store whole 128-byte lines, four rows per
            // instruction.  The activation ring is idle here: this CTA's only item of the phase has been consumed
            // and the next activation tiles arrive after the grid barrier (hence the n_items <= grid condition).
            float* tile = reinterpret_cast<float*>(smem_a + (warp - CH_EPI_WARP0) * (32 * 36 * 4));     // pitch 36 floats: conflict-free float4 rows
#pragma unroll
            for (int k = 0; k < 8; ++k)
              *reinterpret_cast<float4*>(tile + lane * 36 + 4 * k) =
                  make_float4(__uint_as_float(r[4 * k]), __uint_as_float(r[4 * k + 1]), __uint_as_float(r[4 * k + 2]),
                              __uint_as_float(r[4 * k + 3]));
            __syncwarp();
            const size_t zoff = P.part_zs ? static_cast<size_t>(z) * P.part_zs : static_cast<size_t>(z) * g.M * g.ldo;
            float* outp = reinterpret_cast<float*>(g.out) + zoff + col0 + 4 * (lane & 7);
#pragma unroll
            for (int it = 0; it < 8; ++it) {
              const int rr = it * 4 + (lane >> 3);
              const int grow = m_blk * BLOCK_M + q * 32 + rr;
              const float4 v = *reinterpret_cast<const float4*>(tile + rr * 36 + 4 * (lane & 7));
              if (grow < g.M) *reinterpret_cast<float4*>(outp + static_cast<size_t>(grow) * g.ldo) = v;
            }
            if (tid == 0) CTRACE2(ph, 0, 13, clock64());
            fence_proxy_async_smem();          // this staging is the TMA ring again after the barrier
            if (tid == 0) CTRACE2(ph, 0, 14, clock64());
          } else if (row < g.M) {
            if (g.mode == GM_SWIGLU_BF16 && col0 + 32 <= g.N) {
              // 16-byte stores: the lanes of a warp own different rows, so every store instruction is 32
              // separate sector writes whatever its width - halve their number
#pragma unroll
              for (int k = 0; k < 2; ++k) {
                uint32_t w[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  const uint32_t* a = r + k * 16 + i * 4;
                  w[i] = pack_bf16x2(silu_f(__uint_as_float(a[0])) * __uint_as_float(a[1]),
                                     silu_f(__uint_as_float(a[2])) * __uint_as_float(a[3]));
                }
                if (tid == 0) CTRACE2(ph, 0, 13 + k, clock64());
                *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(g.out) + static_cast<size_t>(row) * g.ldo +
                                          ((col0 + k * 16) >> 1)) = make_uint4(w[0], w[1], w[2], w[3]);
              }
              if (tid == 0) CTRACE2(ph, 0, 15, clock64());
            } else {
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const int col = col0 + k * 8;
                if (col < g.N) {
                  if (g.mode == GM_PARTIAL_F32 && P.part_zs) {      // row-major partials: shift the base, store as split 0
                    GemmKParams g2 = g;
                    g2.out = reinterpret_cast<float*>(g.out) + static_cast<size_t>(z) * P.part_zs;
                    epilogue8(g2, row, col, 0, r + k * 8);
                  } else {
                    epilogue8(g, row, col, z, r + k * 8);
                  }
                }
              }
            }
          }
        }
      } else if constexpr (STAGE) {
        if (P.op == CH_DWMIX) phase_dwmix(P.dw, call.cv, rc);
        else phase_dwtile(P.dt, call.cv, rc);
      } else {
        if (P.op == CH_NORM_MOD) {
          phase_norm_mod(P.nm, rc, ph);
        } else if (P.op == CH_TAIL) {
          if (P.tl.split) phase_tail_split(P.tl, call, rc, wr, ph, epoch_base + static_cast<uint32_t>(ph) + 1u);
          else phase_tail(P.tl, call, rc, wr, ph);
        } else {
          phase_cfg(P.cf, call, tid);
        }
      }
      if (tid == 0) CTRACE(16 * ph, clock64());
      grid_sync<false>(flags, epoch_base + static_cast<uint32_t>(ph) + 1u, ph, push);
      if (tid == 0) { CTRACE(16 * ph + 4, clock64()); CTRACE(16 * ph + 7, gtimer()); }
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, CH_TMEM_COLS);
  }
}

}  // namespace

// ------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------
struct ChainProgram {
  std::vector<ChainPhase> phases;
  std::vector<CUtensorMap> tmaps;
  DevBuf d_phases, d_lite, d_tmaps, d_ctr, d_mail;
  uint32_t epoch_base = 0;
  int grid = 0, stages = 0, wstages = 0, smem_bytes = 0;
  long long algo_bytes = 0;      // algorithmic bytes of one launch (bench.py's roofline)
  int rows = 0, gemm_phases = 0;
  int kind = 1;                  // timing-record kind (1: the head's solver loop; 2: a decoder / encoder stage)
  bool finalized = false;
  bool stage_kernel = false;     // k_chain<true> (mixer phases) or k_chain<false> (the head's row phases)
};

ChainProgram* chain_create() { return new ChainProgram(); }
void chain_destroy(ChainProgram* p) { delete p; }
void chain_clear(ChainProgram* p) {
  p->phases.clear();
  p->tmaps.clear();
  p->finalized = false;
  p->algo_bytes = 0;
  p->rows = 0;
  p->gemm_phases = 0;
}
int chain_num_phases(const ChainProgram* p) { return static_cast<int>(p->phases.size()); }

int chain_add_gemm(ChainProgram* p, const GemmArgs& a, int part_rowmajor) {
  VV_REQUIRE(a.A && a.B && a.out, "chain gemm: null operand");
  VV_REQUIRE(a.M >= 1, "chain gemm: M=%d", a.M);
  VV_REQUIRE(a.N % 8 == 0 && a.K % 8 == 0 && a.N > 0 && a.K > 0, "chain gemm: N=%d K=%d", a.N, a.K);
  VV_REQUIRE(a.mode == GM_STORE_F32 || a.mode == GM_SWIGLU_BF16 || a.mode == GM_PARTIAL_F32 ||
                 a.mode == GM_STORE_BF16 || a.mode == GM_SILU_BF16 || a.mode == GM_GELU_BF16 ||
                 (a.mode == GM_RESID_GAMMA_BF16 && a.resid && a.gamma && a.ldr % 8 == 0),
             "chain gemm: mode %d not supported in a chain", a.mode);
  const int splits = a.splits < 1 ? 1 : a.splits;
  VV_REQUIRE(splits == 1 || a.mode == GM_PARTIAL_F32, "chain gemm: only GM_PARTIAL_F32 may split K");
  ChainPhase ph{};
  ph.op = CH_GEMM;
  CUtensorMap tA, tB;
  const int a_rows = a.M <= 64 ? (a.M + 7) / 8 * 8 : BLOCK_M;     // transfer only the rows that exist (see vv_gemm.cu)
  VV_TRY(gemm_make_tmap(&tA, a.A, a.M, a.K, a.lda, a_rows));
  VV_TRY(gemm_make_tmap(&tB, a.B, a.N, a.K, a.ldb, CH_BN));
  ph.tmA = static_cast<int>(p->tmaps.size());
  p->tmaps.push_back(tA);
  ph.tmB = static_cast<int>(p->tmaps.size());
  p->tmaps.push_back(tB);
  ph.n_tiles_n = vv_ceil_div(a.N, CH_BN);
  ph.n_tiles_m = vv_ceil_div(a.M, BLOCK_M);
  ph.n_items = ph.n_tiles_n * ph.n_tiles_m * splits;
  ph.part_zs = 0;
  if (part_rowmajor && a.mode == GM_PARTIAL_F32) {
    VV_REQUIRE(a.ldo == a.N, "chain gemm: row-major partials need ldo == N (ldo=%d N=%d)", a.ldo, a.N);
    ph.part_zs = a.N;
  }
  GemmKParams& g = ph.g;
  g.M = a.M; g.N = a.N; g.K = a.K; g.mode = a.mode;
  g.out = a.out; g.ldo = ph.part_zs ? splits * a.N : a.ldo; g.bias = a.bias; g.gamma = a.gamma;
  g.resid = a.resid; g.ldr = a.ldr; g.gate = a.gate; g.ldg = a.ldg;
  const int kb_total = vv_ceil_div(a.K, BLOCK_K);
  g.kb_per_split = vv_ceil_div(kb_total, splits);
  VV_REQUIRE(g.kb_per_split * (splits - 1) < kb_total, "chain gemm: a split would be empty (K=%d splits=%d)", a.K, splits);
  g.stages = CH_MAX_A_STAGES;
  // weights are read once per phase: evict-first, unless the caller wants this matrix kept in L2 across phases
  // (b_streaming == 2: the head re-reads the same weights every solver step)
  g.a_policy = VV_L2_EVICT_NORMAL; g.b_policy = (a.b_streaming == 2) ? VV_L2_EVICT_LAST : VV_L2_EVICT_FIRST;
  g.a_rows = a_rows;
  {
    // roofline accounting: compulsory HBM bytes only - every weight once per phase that uses it.  Activations,
    // split-K partials and the solver state are scratch that lives in the 126 MB L2 between phases.
    p->algo_bytes += 2ll * a.N * a.K;
    p->rows = a.M;
    ++p->gemm_phases;
  }
  p->phases.push_back(ph);
  p->finalized = false;
  return VV_OK;
}

int chain_add_norm_mod(ChainProgram* p, const ChainNormMod& a) {
  VV_REQUIRE(a.x && a.y && a.shift && a.scale && a.H % 8 == 0 && a.H <= 4096 && a.M >= 1, "chain norm_mod: bad arguments");
  VV_REQUIRE(a.ld_mod % 8 == 0, "chain norm_mod: modulation rows must be 16-byte aligned (ld=%d)", a.ld_mod);
  ChainPhase ph{};
  ph.op = CH_NORM_MOD;
  ph.nm = a;
  p->algo_bytes += static_cast<long long>(a.M) * a.H * 2 * (2 + (a.nsplit > 0 ? 1 : 0));   // shift, scale(, gate) rows of the adaLN table
  p->phases.push_back(ph);
  p->finalized = false;
  return VV_OK;
}

int chain_add_cfg(ChainProgram* p, const ChainCfg& a) {
  VV_REQUIRE(a.v && a.zstate && a.prev_x0 && a.zb && a.n >= 1, "chain cfg: bad arguments");
  ChainPhase ph{};
  ph.op = CH_CFG;
  ph.cf = a;
  p->phases.push_back(ph);
  p->finalized = false;
  return VV_OK;
}

int chain_add_tail(ChainProgram* p, const ChainTail& a) {
  VV_REQUIRE(a.x && a.fshift && a.w_out && a.zstate && a.prev_x0 && a.zb && a.n >= 1, "chain tail: bad arguments");
  VV_REQUIRE(a.last || (a.w_in && a.norm_w0 && a.shift0 && a.y), "chain tail: bad arguments (next step)");
  VV_REQUIRE(a.L == 64 && a.H % 128 == 0 && a.H <= 4096, "chain tail: L=%d H=%d not supported", a.L, a.H);
  VV_REQUIRE(a.nsplit == 0 || (a.partial && a.prev_gate), "chain tail: partials without gate");
  ChainPhase ph{};
  ph.op = CH_TAIL;
  CUtensorMap tO, tI;                                  // w_out [L, H] and w_in [H, L] as [64 x 64] tiles
  VV_TRY(gemm_make_tmap(&tO, a.w_out, a.L, a.H, a.H, 64));
  VV_TRY(gemm_make_tmap(&tI, a.last ? a.w_out : a.w_in, a.last ? a.L : a.H, a.last ? a.H : a.L, a.last ? a.H : a.L, 64));
  ph.tmA = static_cast<int>(p->tmaps.size());
  p->tmaps.push_back(tO);
  ph.tmB = static_cast<int>(p->tmaps.size());
  p->tmaps.push_back(tI);
  ph.tl = a;
  ph.tl.split = 0;                                     // decided in chain_finalize (needs the grid)
  ph.tl.mail = nullptr;
  p->algo_bytes += 2ll * a.L * a.H + 2ll * a.n * a.H * 2 * 3                                 // w_out; gate, fshift, fscale rows
                   + (a.last ? 0ll : 2ll * a.L * a.H + 2ll * a.n * a.H * 2 * 2);                 // w_in; shift0, scale0 rows
  p->phases.push_back(ph);
  p->finalized = false;
  return VV_OK;
}

int chain_add_dwmix(ChainProgram* p, const ChainDw& a) {
  VV_REQUIRE(a.x_out && a.y && a.norm_w && a.dw_w && a.dw_b && a.gamma && a.ffn_norm_w, "chain dwmix: null argument");
  VV_REQUIRE((a.part == nullptr) ? (a.x_in != nullptr) : (a.xres && a.gamma2 && a.bias2 && a.nsplit >= 1),
             "chain dwmix: needs x_in, or partials with their residual");
  VV_REQUIRE(a.rows >= 1 && a.C % 8 == 0 && a.C >= 8 && a.C <= 8 * CH_EPI_THREADS, "chain dwmix: rows=%d C=%d", a.rows, a.C);
  ChainPhase ph{};
  ph.op = CH_DWMIX;
  ph.dw = a;
  p->phases.push_back(ph);
  p->finalized = false;
  return VV_OK;
}
int chain_add_dwtile(ChainProgram* p, const ChainDwTile& a) {
  VV_REQUIRE(a.x_in && a.x_out && a.y && a.norm_w && a.dw_w && a.dw_b && a.gamma && a.ffn_norm_w, "chain dwtile: null argument");
  VV_REQUIRE(a.n >= 1 && a.Tn >= 1 && a.TT >= 1 && a.C % 8 == 0 && a.C >= 8 && a.C <= 8 * CH_EPI_THREADS &&
                 CH_EPI_THREADS % (a.C / 8) == 0,
             "chain dwtile: n=%d Tn=%d TT=%d C=%d", a.n, a.Tn, a.TT, a.C);
  ChainPhase ph{};
  ph.op = CH_DWTILE;
  ph.dt = a;
  p->phases.push_back(ph);
  p->finalized = false;
  return VV_OK;
}
void chain_set_kind(ChainProgram* p, int kind) { p->kind = kind; }

int chain_finalize(ChainProgram* p) {
  VV_REQUIRE(!p->phases.empty(), "chain: empty program");
  static bool attr_flags[VV_MAX_DEVICES] = {};
  bool& attr_set = attr_flags[vv_current_device()];   // function attributes are per device
  if (!attr_set) {
    VV_CHECK_CUDA(cudaFuncSetAttribute(k_chain<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_SMEM));
    VV_CHECK_CUDA(cudaFuncSetAttribute(k_chain<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_SMEM));
    attr_set = true;
  }
  {                                                   // which instantiation walks this program
    bool stage_ops = false, head_ops = false;
    for (const ChainPhase& ph : p->phases) {
      if (ph.op == CH_DWMIX || ph.op == CH_DWTILE) stage_ops = true;
      if (ph.op == CH_NORM_MOD || ph.op == CH_TAIL || ph.op == CH_CFG) head_ops = true;
    }
    VV_REQUIRE(!(stage_ops && head_ops), "chain: a program cannot mix stage (mixer) phases with head row phases");
    p->stage_kernel = stage_ops;
  }
  // Ring sizing: the activation ring doubles as the row phases' staging area, so it must hold the
  // largest staged row set; the weight ring takes what is left (it is what keeps HBM busy across barriers).
  // Both depths are even: a ring slot must always belong to the same one of the two MMA issuers (k_chain).
  int need = 0;
  for (const ChainPhase& ph : p->phases) {
    int nd = 0;
    if (ph.op == CH_NORM_MOD) nd = row_stage_bytes(ph.nm.H, ph.nm.nsplit) + 4 * ph.nm.H;
    if (ph.op == CH_TAIL)                            // one whole row staged, or (split) both rows' halves at once
      nd = std::max(tail_scratch_bytes(ph.tl.H, ph.tl.nsplit),
                    CH_TAIL_FIXED(ph.tl.H) + 2 * ((row_stage_bytes(ph.tl.H / 2, ph.tl.nsplit) + 127) / 128 * 128));
    if (ph.op == CH_DWTILE) nd = (ph.dt.TT + HIST) * ph.dt.C * 2;
    need = std::max(need, nd);
  }
  const int avail = MAX_SMEM - 1024 - CH_CTRL_BYTES;
  int sa = std::max({CH_MIN_A_STAGES, vv_ceil_div(need, CH_A_BYTES), std::min(vv_option_chain_stages(), CH_MAX_A_STAGES)});
  sa += sa & 1;
  VV_REQUIRE(sa <= CH_MAX_A_STAGES, "chain: a row phase needs %d bytes of staging, more than %d activation stages", need, CH_MAX_A_STAGES);
  int sb = (avail - sa * CH_A_BYTES) / CH_B_BYTES;
  sb = std::min({sb, CH_MAX_W_STAGES, std::max(vv_option_chain_wstages(), CH_MIN_W_STAGES)});
  sb -= sb & 1;
  VV_REQUIRE(sb >= CH_MIN_W_STAGES, "chain: program of %d phases leaves no room for the weight ring", (int)p->phases.size());
  p->stages = sa;
  p->wstages = sb;
  p->smem_bytes = 1024 + sa * CH_A_BYTES + sb * CH_B_BYTES + CH_CTRL_BYTES;
  int per_sm = 0;
  if (p->stage_kernel) VV_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_chain<true>, CH_THREADS, p->smem_bytes));
  else VV_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_chain<false>, CH_THREADS, p->smem_bytes));
  VV_REQUIRE(per_sm >= 1, "chain: kernel does not fit on an SM");
  p->grid = vv_num_sms();
  // step-boundary phases: two CTAs per row pair where the shape allows it (ChainTail::split)
  {
    int max_n = 0;
    for (ChainPhase& ph : p->phases)
      if (ph.op == CH_TAIL) {
        ph.tl.split = (vv_option_head_tail_split() && ph.tl.H % 256 == 0 && p->grid >= 2) ? 1 : 0;
        if (ph.tl.split) max_n = std::max(max_n, ph.tl.n);
      }
    if (max_n > 0) {
      const size_t bytes = static_cast<size_t>(max_n) * 2 * CH_MAIL_FLOATS * sizeof(float);
      if (p->d_mail.bytes < bytes) {
        VV_TRY(p->d_mail.alloc(bytes));
        VV_CHECK_CUDA(cudaMemset(p->d_mail.p, 0, bytes));
      }
      for (ChainPhase& ph : p->phases)
        if (ph.op == CH_TAIL && ph.tl.split) ph.tl.mail = p->d_mail.as<float>();
    }
  }
  // the 16-byte view of every phase that the producers and the MMA issuers read (ChainLite)
  std::vector<ChainLite> lite(p->phases.size());
  for (size_t i = 0; i < p->phases.size(); ++i) {
    const ChainPhase& ph = p->phases[i];
    ChainLite& l = lite[i];
    memset(&l, 0, sizeof(l));
    l.op = static_cast<uint8_t>(ph.op);
    l.tmA = static_cast<int16_t>(ph.tmA);
    l.tmB = static_cast<int16_t>(ph.tmB);
    if (ph.op == CH_GEMM) {
      const int kb_total = (ph.g.K + BLOCK_K - 1) / BLOCK_K;
      VV_REQUIRE(ph.n_items <= 65535 && ph.n_tiles_n <= 65535 && ph.n_tiles_m <= 255 && kb_total <= 65535 &&
                     ph.g.kb_per_split <= 65535 && ph.g.a_rows <= 255 && ph.tmA <= 32767 && ph.tmB <= 32767,
                 "chain: GEMM phase %d exceeds the compact phase table's ranges", (int)i);
      l.flag = ph.g.b_policy == VV_L2_EVICT_LAST ? 1 : 0;
      l.n_tiles_m = static_cast<uint8_t>(ph.n_tiles_m);
      l.a_rows = static_cast<uint8_t>(ph.g.a_rows);
      l.n_items = static_cast<uint16_t>(ph.n_items);
      l.n_tiles_n = static_cast<uint16_t>(ph.n_tiles_n);
      l.kb_total = static_cast<uint16_t>(kb_total);
      l.kb_per_split = static_cast<uint16_t>(ph.g.kb_per_split);
    } else if (ph.op == CH_TAIL) {
      VV_REQUIRE(ph.tl.n <= 65535 && ph.tl.H % 64 == 0, "chain: step-boundary phase %d: n=%d H=%d", (int)i, ph.tl.n, ph.tl.H);
      l.flag = ph.tl.last ? 1 : 0;
      l.n_tiles_m = static_cast<uint8_t>(ph.tl.split);
      l.n_items = static_cast<uint16_t>(ph.tl.n);
      l.kb_total = static_cast<uint16_t>(ph.tl.H / 64);
    }
  }
  VV_TRY(p->d_lite.alloc(lite.size() * sizeof(ChainLite)));
  VV_TRY(p->d_phases.alloc(p->phases.size() * sizeof(ChainPhase)));
  VV_TRY(p->d_tmaps.alloc(p->tmaps.size() * sizeof(CUtensorMap)));
  if (!p->d_ctr.p) {
    // one 128-byte counter line per CTA (grid_sync), behind an unused first line
    VV_TRY(p->d_ctr.alloc(128 * (1 + CH_MAX_GRID)));
    VV_CHECK_CUDA(cudaMemset(p->d_ctr.p, 0, 128 * (1 + CH_MAX_GRID)));
    p->epoch_base = 0;
  }
  VV_COUNT_SYNC();                                          // synchronous uploads below
  VV_CHECK_CUDA(cudaMemcpy(p->d_phases.p, p->phases.data(), p->phases.size() * sizeof(ChainPhase), cudaMemcpyHostToDevice));
  VV_CHECK_CUDA(cudaMemcpy(p->d_lite.p, lite.data(), lite.size() * sizeof(ChainLite), cudaMemcpyHostToDevice));
  VV_CHECK_CUDA(cudaMemcpy(p->d_tmaps.p, p->tmaps.data(), p->tmaps.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice));
  p->finalized = true;
  return VV_OK;
}

int chain_launch(ChainProgram* p, const ChainCall& call, cudaStream_t stream) {
  VV_REQUIRE(p->finalized, "chain: program not finalized");
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(p->grid);
  cfg.blockDim = dim3(CH_THREADS);
  cfg.dynamicSmemBytes = p->smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;      // all CTAs co-resident: the grid barrier cannot deadlock
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const int n_phases = static_cast<int>(p->phases.size());
  const int push = vv_option_chain_push_barrier();   // which cross-proxy fences the grid barrier uses (grid_sync)
  VV_COUNT_LAUNCH();
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  const bool timed = gemm_timing_on();
  if (timed) {
    VV_CHECK_CUDA(cudaEventCreate(&e0));
    VV_CHECK_CUDA(cudaEventCreate(&e1));
    VV_CHECK_CUDA(cudaEventRecord(e0, stream));
  }
  VV_CHECK_CUDA(cudaLaunchKernelEx(&cfg, p->stage_kernel ? k_chain<true> : k_chain<false>, static_cast<const ChainPhase*>(p->d_phases.as<ChainPhase>()),
                                   static_cast<const ChainLite*>(p->d_lite.as<ChainLite>()), n_phases, p->stages,
                                   p->wstages, static_cast<const CUtensorMap*>(p->d_tmaps.as<CUtensorMap>()), call,
                                   p->d_ctr.as<uint32_t>(), p->epoch_base, vv_option_chain_krot(), push));
  p->epoch_base += static_cast<uint32_t>(n_phases);
  if (timed) {
    VV_CHECK_CUDA(cudaEventRecord(e1, stream));
    gemm_timing_push(GemmLaunchRec{p->rows, n_phases, p->gemm_phases, CH_BN, 1, -1, 0.f, p->kind, p->algo_bytes}, e0, e1);
  }
  return VV_OK;
}

int chain_debug_trace(long long* out, int cap) {
#ifdef VV_GEMM_TRACE
  const int n = 160 * 384, n2 = 160 * 24 * 3 * 16;
  if (cap < n) return -1;
  if (cudaMemcpyFromSymbol(out, g_chain_trace, sizeof(long long) * n) != cudaSuccess) return -2;
  if (cap >= n + n2) {                          // the per-group stamps follow when the caller left room for them
    if (cudaMemcpyFromSymbol(out + n, g_chain_trace2, sizeof(long long) * n2) != cudaSuccess) return -2;
    return n + n2;
  }
  return n;
#else
  (void)out; (void)cap;
  return 0;
#endif
}
