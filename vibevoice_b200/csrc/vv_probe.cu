// vv_probe.cu — measurement aid, not on the product path: how fast can one CTA per SM pull a weight
// stream through a TMA -> mbarrier ring when nothing consumes it?  Answers "what is the ceiling of the
// weight-streaming GEMMs" as a function of box size, ring depth, CTAs per SM and copy flavour
// (profiles/r01_tma_stream.md).  Each CTA streams its own contiguous slab of a [rows, 64] bf16 tensor
// (128-byte rows), `box_rows` rows per box; a second thread releases every slot as soon as it lands.
#include "../../include/vvb200_probe.h"
#include "vv_gemm.cuh"

namespace {
constexpr int PROBE_MAX_STAGES = 32;

__global__ void __launch_bounds__(64, 1)
k_tma_probe(const __grid_constant__ CUtensorMap tm, const __grid_constant__ CUtensorMap tm2, const uint8_t* __restrict__ base, int boxes_per_cta,
            int box_rows, int stages, int flavour, int shared_slab) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int box_bytes = box_rows * 128;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + stages * box_bytes);
  uint64_t* empty_bar = full_bar + PROBE_MAX_STAGES;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    fence_mbar_init();
    tma_prefetch_desc(&tm);
  }
  __syncthreads();
  // shared_slab: every CTA reads the SAME boxes (what the activation tiles of a skinny GEMM do)
  const long long first = shared_slab ? 0 : static_cast<long long>(blockIdx.x) * boxes_per_cta;
  if (threadIdx.x == 0) {
    int stage = 0;
    uint32_t phase = 0;
    for (int i = 0; i < boxes_per_cta; ++i) {
      mbar_wait(&empty_bar[stage], phase ^ 1);
      uint8_t* dst = smem + stage * box_bytes;
      if (flavour == 2) {
        // the skinny GEMM's pattern: per k-block one private 8 KB weight box (HBM stream) plus one 16 KB
        // activation box that every CTA shares (L2); box_rows must be 192 (24 KB per stage)
        mbar_arrive_expect_tx(&full_bar[stage], 24 * 1024);
        tma_load_2d(dst, &tm, &full_bar[stage], 0, (i & 63) * 128, VV_L2_EVICT_NORMAL);                     // shared, 128 rows (box = 128)
        tma_load_2d(dst + 16 * 1024, &tm2, &full_bar[stage], 0,
                    static_cast<int>(8192 + (static_cast<long long>(blockIdx.x) * boxes_per_cta + i) * 64),
                    VV_L2_EVICT_FIRST);                                                                      // private, 64 rows (box = 64)
      } else {
        mbar_arrive_expect_tx(&full_bar[stage], box_bytes);
        if (flavour == 0) {
          tma_load_2d(dst, &tm, &full_bar[stage], 0, static_cast<int>((first + i) * box_rows), VV_L2_EVICT_FIRST);
        } else {
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                       ::"r"(smem_u32(dst)), "l"(base + (first + i) * box_bytes), "r"(box_bytes),
                         "r"(smem_u32(&full_bar[stage])) : "memory");
        }
      }
      if (++stage == stages) { stage = 0; phase ^= 1; }
    }
  } else if (threadIdx.x == 32) {
    int stage = 0;
    uint32_t phase = 0;
    for (int i = 0; i < boxes_per_cta; ++i) {
      mbar_wait(&full_bar[stage], phase);
      mbar_arrive(&empty_bar[stage]);
      if (++stage == stages) { stage = 0; phase ^= 1; }
    }
  }
}
// What does a tcgen05.commit cost the issuing thread?  One thread issues `mmas` 128 x bn x 16 MMAs on garbage
// operands already in shared memory, then `commits` commits, `iters` times; out[cta] = clk per iteration.
__global__ void __launch_bounds__(64, 1) k_mma_commit_probe(int bn, int mmas, int commits, int iters, int nacc, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t bars[6];
  __shared__ uint32_t tmem_slot;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 6; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
  }
  if (threadIdx.x < 32) tmem_alloc(&tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t d = tmem_slot;
  if (nacc <= -2) {
    // nacc = -2 / -3: TWO converged warps issue `mmas` MMAs + `commits` commits per iteration, each into its own
    // accumulator (and with two waits on already-complete mbarriers per iteration when nacc = -3, the shape of the
    // phase engine's main loop); nacc = -4: one warp with the two waits.  Does a second issuing warp double the rate?
    const int w = threadIdx.x >> 5;
    if (w == 0 || (w == 1 && nacc != -4)) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(bn >> 3) << 17) | (static_cast<uint32_t>(128 >> 4) << 24);
      const uint64_t adesc = umma_desc_sw128(smem_u32(smem + w * 8192));
      const uint64_t bdesc = umma_desc_sw128(smem_u32(smem + w * 8192 + 16384));
      const uint32_t dd = d + w * 128;
      const long long t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        if (nacc != -2) {
          mbar_wait(&bars[3], 1);          // never-armed barrier: waiting for the phase before its first completes at once
          mbar_wait(&bars[3], 1);
          tc_fence_after();
        }
        for (int g = 0; g < mmas; g += 4) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16_warp(dd, adesc + 2 * k, bdesc + 2 * k, idesc, 1u);
        }
        for (int c = 0; c < commits; ++c) umma_commit_warp(&bars[w]);
      }
      umma_commit_warp(&bars[4 + w]);
      mbar_wait(&bars[4 + w], 0);
      const long long t1 = clock64();
      if (threadIdx.x == 0) out[blockIdx.x] = (t1 - t0) / iters;
    }
  } else if (nacc < 0 && threadIdx.x < 32) {        // converged warp, elect.sync-predicated issue
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(bn >> 3) << 17) | (static_cast<uint32_t>(128 >> 4) << 24);
    const uint64_t adesc = umma_desc_sw128(smem_u32(smem));
    const uint64_t bdesc = umma_desc_sw128(smem_u32(smem + 16384));
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      for (int g = 0; g < mmas; g += 4) {
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16_warp(d, adesc + 2 * k, bdesc + 2 * k, idesc, 1u);
      }
      for (int c = 0; c < commits; ++c) umma_commit_warp(&bars[c & 1]);
    }
    umma_commit_warp(&bars[2]);
    mbar_wait(&bars[2], 0);
    const long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = (t1 - t0) / iters;
  } else if (nacc >= 0 && threadIdx.x == 0) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(bn >> 3) << 17) | (static_cast<uint32_t>(128 >> 4) << 24);
    const uint64_t adesc = umma_desc_sw128(smem_u32(smem));
    const uint64_t bdesc = umma_desc_sw128(smem_u32(smem + 16384));
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      if (nacc == 0) {                    // four MMAs, fully unrolled, one accumulator (what the GEMM main loops do)
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16(d, adesc + 2 * k, bdesc + 2 * k, idesc, 1u);
      } else {
        for (int k = 0; k < mmas; ++k) umma_bf16(d + (k % nacc) * bn, adesc + 2 * (k & 3), bdesc + 2 * (k & 3), idesc, 1u);
      }
      for (int c = 0; c < commits; ++c) umma_commit(&bars[c & 1]);
    }
    umma_commit(&bars[2]);
    mbar_wait(&bars[2], 0);
    const long long t1 = clock64();
    out[blockIdx.x] = (t1 - t0) / iters;
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    tc_fence_after();
    tmem_dealloc(d, 256);
  }
}
}  // namespace

extern "C" int vv_debug_mma_commit(int bn, int mmas, int commits, int iters, long long* out_host, int ctas, int nacc) {
  VV_REQUIRE(bn >= 16 && bn <= 256 && bn % 16 == 0 && mmas >= 1 && commits >= 0 && iters >= 1 && ctas >= 1 && ctas <= 1024, "probe: bad arguments");
  DevBuf o;
  VV_TRY(o.alloc(sizeof(long long) * ctas));
  VV_CHECK_CUDA(cudaFuncSetAttribute(k_mma_commit_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  VV_REQUIRE(nacc >= -4 && nacc * bn <= 256, "probe: nacc");
  k_mma_commit_probe<<<ctas, 64, 56 * 1024>>>(bn, mmas, commits, iters, nacc, o.as<long long>());
  VV_CHECK_CUDA(cudaGetLastError());
  VV_CHECK_CUDA(cudaDeviceSynchronize());
  VV_CHECK_CUDA(cudaMemcpy(out_host, o.p, sizeof(long long) * ctas, cudaMemcpyDeviceToHost));
  return VV_OK;
}

extern "C" int vv_debug_tma_stream(const void* buf, long long bytes, int box_rows, int stages, int ctas, int flavour,
                                   int shared_slab, void* stream) {
  VV_REQUIRE(buf && box_rows >= 8 && box_rows <= 256 && box_rows % 8 == 0, "probe: box_rows=%d", box_rows);
  if (flavour == 2) bytes = bytes / 4;   // only the 8 KB of each 24 KB stage is private data: keep the slabs inside the buffer
  VV_REQUIRE(stages >= 1 && stages <= PROBE_MAX_STAGES && ctas >= 1, "probe: stages=%d ctas=%d", stages, ctas);
  const int box_bytes = box_rows * 128;
  const int smem = 1024 + stages * box_bytes + 2 * PROBE_MAX_STAGES * 8;
  VV_REQUIRE(smem <= 227 * 1024, "probe: ring of %d bytes does not fit", smem);
  const long long boxes = bytes / box_bytes;
  const int per_cta = static_cast<int>(shared_slab ? boxes : boxes / ctas);
  VV_REQUIRE(per_cta >= 1, "probe: buffer too small");
  CUtensorMap tm;
  const long long rows = boxes * box_rows;
  VV_REQUIRE(rows < (1ll << 31), "probe: buffer too large");
  CUtensorMap tm2;
  if (flavour == 2) {
    VV_REQUIRE(box_rows == 192 && !shared_slab, "probe: flavour 2 needs box_rows = 192, private slabs");
    VV_TRY(gemm_make_tmap(&tm, static_cast<const bf16*>(buf), static_cast<int>(rows), 64, 64, 128));
    VV_TRY(gemm_make_tmap(&tm2, static_cast<const bf16*>(buf), static_cast<int>(rows), 64, 64, 64));
  } else {
    VV_TRY(gemm_make_tmap(&tm, static_cast<const bf16*>(buf), static_cast<int>(rows), 64, 64, box_rows));
    tm2 = tm;
  }
  VV_CHECK_CUDA(cudaFuncSetAttribute(k_tma_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  k_tma_probe<<<ctas, 64, smem, static_cast<cudaStream_t>(stream)>>>(tm, tm2, static_cast<const uint8_t*>(buf), per_cta,
                                                                      box_rows, stages, flavour, shared_slab);
  VV_CHECK_CUDA(cudaGetLastError());
  return VV_OK;
}
