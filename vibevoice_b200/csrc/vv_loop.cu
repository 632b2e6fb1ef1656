// vv_loop.cu — device-side bookkeeping of the frame loop.
//
// generate() (modeling_vibevoice_inference.py:556-670) decides on the HOST, every frame, which rows are in diffusion
// mode: `torch.arange(batch)[~finished & (next_tokens == speech_diffusion_id)]` (:588, a nonzero => one sync),
// `if diffusion_indices.numel() > 0`, then per active row `sample_idx.item()` + `finished_tags[idx]` + a Python
// list append (:653-658), and the streamer's per-row `.item()` / `.cpu()` (streamer.py:42-55).  At ~2.6 ms of GPU
// time per frame those syncs are the frame loop's bound.  Here the row set never leaves the device:
//
//   * vv_frame_select   mask -> slot map: an active row keeps its own cache slot, an inactive row is pointed at a
//                       spare "trash" slot.  The whole batch then runs through head / decode / encode / connectors
//                       at a FIXED launch shape; what an inactive row computes lands in the trash slot or is dropped
//                       by the masked writes below.  Also emits the ordered compaction (the reference's
//                       diffusion_indices) and the count, on the device, for callers that want them.
//   * vv_ring_append    audio_chunks[idx].append(chunk[i]) for every active row at once: chunk -> ring[row, pos[row]],
//                       pos[row] += 1.  One D2H at the end of generation replaces torch.cat over Python lists (:694-700).
//   * vv_embed_scatter  next_inputs_embeds[diffusion_indices] = diffusion_embeds (:687) as a masked row copy.
#include <algorithm>

#include "../../include/vvb200.h"
#include "vv_common.cuh"

namespace {

// One block.  active[i] = !finished[i] && tokens[i] == token_id (finished == nullptr: no row is finished;
// tokens == nullptr: `finished` is read as the mask itself, inverted: active = !finished).
__global__ void __launch_bounds__(1024) k_frame_select(const long long* __restrict__ tokens,
                                                       const unsigned char* __restrict__ finished, int n,
                                                       long long token_id, long long trash, long long* __restrict__ slots,
                                                       unsigned char* __restrict__ active, long long* __restrict__ compact,
                                                       int* __restrict__ count) {
  griddep_wait();
  griddep_launch();
  __shared__ int warp_tot[32];
  __shared__ int base_s;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) base_s = 0;
  __syncthreads();
  for (int i0 = 0; i0 < n; i0 += 1024) {
    const int i = i0 + tid;
    bool a = false;
    if (i < n) {
      const bool fin = finished != nullptr && finished[i] != 0;
      a = !fin && (tokens == nullptr || tokens[i] == token_id);
      if (slots) slots[i] = a ? static_cast<long long>(i) : trash;
      if (active) active[i] = a ? 1 : 0;
    }
    const unsigned bal = __ballot_sync(0xffffffffu, a);
    if (lane == 0) warp_tot[warp] = __popc(bal);
    __syncthreads();
    int before = base_s;
    for (int w = 0; w < warp; ++w) before += warp_tot[w];
    if (a && compact) compact[before + __popc(bal & ((1u << lane) - 1u))] = i;
    __syncthreads();
    if (tid == 0) {
      int t = 0;
      for (int w = 0; w < 32; ++w) t += warp_tot[w];
      base_s += t;
    }
    __syncthreads();
  }
  const int total = base_s;
  if (compact)
    for (int i = total + tid; i < n; i += 1024) compact[i] = -1;
  if (tid == 0 && count) *count = total;
}

template <typename U>
__global__ void k_ring_append(const U* __restrict__ chunk, const unsigned char* __restrict__ active,
                              const int* __restrict__ pos, U* __restrict__ ring, int n, int upr, int cap_frames) {
  griddep_wait();
  griddep_launch();
  const size_t total = static_cast<size_t>(n) * upr;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int b = static_cast<int>(i / upr), u = static_cast<int>(i % upr);
    if (!active[b]) continue;
    const int p = pos[b];
    if (p >= cap_frames) continue;                 // full: the frame is dropped, pos still counts it (caller checks)
    ring[(static_cast<size_t>(b) * cap_frames + p) * upr + u] = chunk[i];
  }
}
__global__ void k_ring_advance(const unsigned char* __restrict__ active, int* __restrict__ pos, int n) {
  griddep_wait();
  griddep_launch();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && active[i]) pos[i] += 1;
}

template <typename T>
__global__ void k_embed_scatter(const bf16* __restrict__ emb, const unsigned char* __restrict__ active,
                                T* __restrict__ dst, int n, int H) {
  griddep_wait();
  griddep_launch();
  const size_t total = static_cast<size_t>(n) * H;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    if (active[i / H]) dst[i] = static_cast<T>(__bfloat162float(emb[i]));
  }
}
template <>
__global__ void k_embed_scatter<bf16>(const bf16* __restrict__ emb, const unsigned char* __restrict__ active,
                                      bf16* __restrict__ dst, int n, int H) {
  griddep_wait();
  griddep_launch();
  const size_t total = static_cast<size_t>(n) * H;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    if (active[i / H]) dst[i] = emb[i];
  }
}

inline unsigned grid_for(size_t total) {
  return static_cast<unsigned>(std::max<size_t>(1, std::min<size_t>((total + 255) / 256, 148u * 8u)));
}

}  // namespace

extern "C" {

int vv_frame_select(const int64_t* next_tokens, const void* finished, int n, int64_t token_id, int64_t trash_slot,
                    int64_t* slots_out, void* active_out, int64_t* compact_out, int* count_out, void* stream) {
  VV_REQUIRE(next_tokens || finished, "vv_frame_select: need next_tokens and/or finished");
  VV_REQUIRE(n >= 1, "vv_frame_select: n=%d", n);
  VV_REQUIRE(slots_out || active_out || compact_out || count_out, "vv_frame_select: no output requested");
  VV_CHECK_CUDA(vv_launch(k_frame_select, dim3(1), dim3(1024), 0, static_cast<cudaStream_t>(stream), 1, true,
                          reinterpret_cast<const long long*>(next_tokens), static_cast<const unsigned char*>(finished), n,
                          static_cast<long long>(token_id), static_cast<long long>(trash_slot),
                          reinterpret_cast<long long*>(slots_out), static_cast<unsigned char*>(active_out),
                          reinterpret_cast<long long*>(compact_out), count_out));
  return VV_OK;
}

int vv_ring_append(const void* chunk, const void* active, int n, int hop, int io_dtype, void* ring, int* pos,
                   int cap_frames, void* stream) {
  VV_REQUIRE(chunk && active && ring && pos, "vv_ring_append: null argument");
  VV_REQUIRE(n >= 1 && hop >= 1 && cap_frames >= 1, "vv_ring_append: n=%d hop=%d cap_frames=%d", n, hop, cap_frames);
  VV_REQUIRE(io_dtype == VV_DTYPE_BF16 || io_dtype == VV_DTYPE_F32, "vv_ring_append: io_dtype");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t row_bytes = static_cast<size_t>(hop) * (io_dtype == VV_DTYPE_BF16 ? 2 : 4);
  const bool v16 = row_bytes % 16 == 0 && reinterpret_cast<uintptr_t>(chunk) % 16 == 0 &&
                   reinterpret_cast<uintptr_t>(ring) % 16 == 0;
  const unsigned char* act = static_cast<const unsigned char*>(active);
  if (v16) {
    const int upr = static_cast<int>(row_bytes / 16);
    VV_CHECK_CUDA(vv_launch(k_ring_append<uint4>, dim3(grid_for(static_cast<size_t>(n) * upr)), dim3(256), 0, st, 1, true,
                            static_cast<const uint4*>(chunk), act, static_cast<const int*>(pos), static_cast<uint4*>(ring),
                            n, upr, cap_frames));
  } else {
    const int upr = static_cast<int>(row_bytes / 2);
    VV_CHECK_CUDA(vv_launch(k_ring_append<unsigned short>, dim3(grid_for(static_cast<size_t>(n) * upr)), dim3(256), 0, st, 1,
                            true, static_cast<const unsigned short*>(chunk), act, static_cast<const int*>(pos),
                            static_cast<unsigned short*>(ring), n, upr, cap_frames));
  }
  VV_CHECK_CUDA(vv_launch(k_ring_advance, dim3((n + 127) / 128), dim3(128), 0, st, 1, true, act, pos, n));
  return VV_OK;
}

int vv_embed_scatter(const void* emb, const void* active, int n, int H, void* dst, int dst_dtype, void* stream) {
  VV_REQUIRE(emb && active && dst, "vv_embed_scatter: null argument");
  VV_REQUIRE(n >= 1 && H >= 1, "vv_embed_scatter: n=%d H=%d", n, H);
  VV_REQUIRE(dst_dtype == VV_DTYPE_BF16 || dst_dtype == VV_DTYPE_F32, "vv_embed_scatter: dst_dtype");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const unsigned g = grid_for(static_cast<size_t>(n) * H);
  if (dst_dtype == VV_DTYPE_BF16)
    VV_CHECK_CUDA(vv_launch(k_embed_scatter<bf16>, dim3(g), dim3(256), 0, st, 1, true, static_cast<const bf16*>(emb),
                            static_cast<const unsigned char*>(active), static_cast<bf16*>(dst), n, H));
  else
    VV_CHECK_CUDA(vv_launch(k_embed_scatter<float>, dim3(g), dim3(256), 0, st, 1, true, static_cast<const bf16*>(emb),
                            static_cast<const unsigned char*>(active), static_cast<float*>(dst), n, H));
  return VV_OK;
}

}  // extern "C"
