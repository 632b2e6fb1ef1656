// vv_api.cu — error plumbing, version, and the generic fused-linear entry point.
#include <stdarg.h>
#include <string.h>

#include <atomic>

#include "../../include/vvb200.h"
#include "vv_common.cuh"
#include "vv_block.cuh"
#include "vv_chain.cuh"
#include "vv_gemm.cuh"

static thread_local char g_err[1024] = "";
std::atomic<long long> g_vv_launches{0};
std::atomic<long long> g_vv_host_syncs{0};
static std::atomic<bool> g_pdl{true};
bool vv_pdl_enabled() { return g_pdl.load(std::memory_order_relaxed); }
static std::atomic<bool> g_fused_block{true};
static std::atomic<bool> g_block_pipe{true};
bool vv_option_block_pipe() { return g_block_pipe.load(std::memory_order_relaxed); }
static std::atomic<int> g_head_chain{1};
int vv_option_head_chain() { return g_head_chain.load(std::memory_order_relaxed); }
static std::atomic<bool> g_gemm_persistent{true};
bool vv_option_gemm_persistent() { return g_gemm_persistent.load(std::memory_order_relaxed); }
static std::atomic<int> g_chain_stages{3};
int vv_option_chain_stages() { return g_chain_stages.load(std::memory_order_relaxed); }
static std::atomic<int> g_gemm_stream_bn32{1};
int vv_option_gemm_stream_bn32() { return g_gemm_stream_bn32.load(std::memory_order_relaxed); }
static std::atomic<int> g_gemm_swap{0};   // measured on B200: no gain (the launches it applies to are bound by fixed per-launch cost)
int vv_option_gemm_swap() { return g_gemm_swap.load(std::memory_order_relaxed); }
static std::atomic<int> g_stage_chain{1};
int vv_option_stage_chain() { return g_stage_chain.load(std::memory_order_relaxed); }
// measured on B200 (n = 64 decoder step): 1.388 ms with the tiled stage programs against 1.305 ms without - 64-wide tiles
// re-read a 128-row activation tile per work item and the mixer tiles are latency-bound either way: off by default
static std::atomic<int> g_stage_tiled{0};
static std::atomic<int> g_dw_rows{0};
int vv_option_dw_rows() { return g_dw_rows.load(std::memory_order_relaxed); }
int vv_option_stage_tiled() { return g_stage_tiled.load(std::memory_order_relaxed); }
static std::atomic<int> g_chain_push{2};
int vv_option_chain_push_barrier() { return g_chain_push.load(std::memory_order_relaxed); }
static std::atomic<int> g_head_tail_split{0};
int vv_option_head_tail_split() { return g_head_tail_split.load(std::memory_order_relaxed); }
static std::atomic<int> g_chain_part_rm{1};
int vv_option_chain_part_rowmajor() { return g_chain_part_rm.load(std::memory_order_relaxed); }
static std::atomic<int> g_head_chain_rows{512};
int vv_option_head_chain_rows() { return g_head_chain_rows.load(std::memory_order_relaxed); }
static std::atomic<int> g_head_fuse_tail{1};
int vv_option_head_fuse_tail() { return g_head_fuse_tail.load(std::memory_order_relaxed); }
static std::atomic<int> g_dw_bulk{1};
int vv_option_dw_bulk() { return g_dw_bulk.load(std::memory_order_relaxed); }
static std::atomic<int> g_head_conv_walk{1};
int vv_option_head_conv_walk() { return g_head_conv_walk.load(std::memory_order_relaxed); }
static std::atomic<int> g_enc_splitk{1};
int vv_option_enc_splitk() { return g_enc_splitk.load(std::memory_order_relaxed); }
static std::atomic<int> g_chain_krot{0};
int vv_option_chain_krot() { return g_chain_krot.load(std::memory_order_relaxed); }
// measured on B200 (n = 64, head alone): off 1.093 ms, interleaved (1) 1.172 ms, one burst behind a phase's last push (2) 1.137 ms -
// the phases do not wait for the weight stream, and the extra TMA requests compete with the loads that are on the critical path
// measured on B200 (head alone, n = 64): 1.099 ms with 0, 1.109 / 1.114 / 1.114 / 1.114 ms with 1..4 layers kept - the GEMM phases are
// bound by the MMA-issue / ring loop, not by where the weight tiles come from
static std::atomic<int> g_chain_keep_layers{0};
int vv_option_chain_keep_layers() { return g_chain_keep_layers.load(std::memory_order_relaxed); }
static std::atomic<int> g_chain_wstages{12};
int vv_option_chain_wstages() { return g_chain_wstages.load(std::memory_order_relaxed); }
static std::atomic<int> g_trace_select{0};   // vv_debug_trace source: 0 GEMM, 1 fused block, 2 head chain, 3 fused FFN
bool vv_option_fused_block() { return g_fused_block.load(std::memory_order_relaxed); }

// ---- launch log (diagnostic) ----
#include <vector>
namespace {
struct LoggedLaunch { const void* fn; cudaEvent_t e0, e1; };
std::atomic<bool> g_launch_log{false};
std::vector<LoggedLaunch> g_logged;
}  // namespace
bool vv_launch_log_on() { return g_launch_log.load(std::memory_order_relaxed); }
void vv_launch_log_begin(const void* fn, cudaStream_t st) {
  LoggedLaunch l{fn, nullptr, nullptr};
  cudaEventCreate(&l.e0);
  cudaEventCreate(&l.e1);
  cudaEventRecord(l.e0, st);
  g_logged.push_back(l);
}
void vv_launch_log_end(cudaStream_t st) {
  if (!g_logged.empty()) cudaEventRecord(g_logged.back().e1, st);
}
static void launch_log_dump() {
  cudaDeviceSynchronize();
  double total = 0;
  for (LoggedLaunch& l : g_logged) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, l.e0, l.e1);
    const char* name = nullptr;
    if (cudaFuncGetName(&name, l.fn) != cudaSuccess || !name) name = "?";
    fprintf(stderr, "vv_launch %8.1f us  %s\n", ms * 1e3, name);
    total += ms;
    cudaEventDestroy(l.e0);
    cudaEventDestroy(l.e1);
  }
  fprintf(stderr, "vv_launch total %.1f us over %zu launches\n", total * 1e3, g_logged.size());
  g_logged.clear();
}

void vv_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int vv_num_sms() {
  static int sms[VV_MAX_DEVICES] = {};
  const int dev = vv_current_device();
  if (sms[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;  // B200
    sms[dev] = v;
  }
  return sms[dev];
}

extern "C" {

int vv_version(void) { return 100; }  // 0.1.0

const char* vv_last_error(void) { return g_err; }

long long vv_launch_count(void) { return g_vv_launches.load(); }

long long vv_host_sync_count(void) { return g_vv_host_syncs.load(); }

int vv_linear_plan(int M, int N, int K, int mode, int* bn, int* splits, int* cluster, int* stages) {
  GemmArgs a;
  a.M = M; a.N = N; a.K = K; a.mode = mode;
  a.bn = bn ? *bn : 0;
  a.splits = splits ? *splits : 0;
  a.cluster = cluster ? *cluster : 0;
  a.stages = stages ? *stages : 0;
  a.swap = 2;
  gemm_plan(a, vv_num_sms());
  if (bn) *bn = a.bn;
  if (splits) *splits = a.splits;
  if (cluster) *cluster = a.cluster;
  if (stages) *stages = a.stages;
  return VV_OK;
}

int vv_set_option(const char* name, int value) {
  if (name && strcmp(name, "pdl") == 0) {
    const int prev = g_pdl.exchange(value != 0) ? 1 : 0;
    return prev;
  }
  if (name && strcmp(name, "launch_log") == 0) {
    const bool prev = g_launch_log.exchange(value != 0);
    if (prev && value == 0) launch_log_dump();
    return prev ? 1 : 0;
  }
  if (name && strcmp(name, "trace_select") == 0) return g_trace_select.exchange(value);
  if (name && strcmp(name, "head_chain") == 0) return g_head_chain.exchange(value < 0 ? 0 : value);
  if (name && strcmp(name, "gemm_persistent") == 0) return g_gemm_persistent.exchange(value != 0) ? 1 : 0;
  if (name && strcmp(name, "chain_stages") == 0) return g_chain_stages.exchange(value < 2 ? 2 : value);
  if (name && strcmp(name, "gemm_stream_bn32") == 0) return g_gemm_stream_bn32.exchange(value != 0);
  if (name && strcmp(name, "gemm_swap") == 0) return g_gemm_swap.exchange(value != 0);
  if (name && strcmp(name, "stage_chain") == 0) return g_stage_chain.exchange(value < 0 ? 0 : value);
  if (name && strcmp(name, "dw_rows") == 0) return g_dw_rows.exchange(value < 0 ? 0 : value);
  if (name && strcmp(name, "stage_tiled") == 0) return g_stage_tiled.exchange(value < 0 ? 0 : value);
  if (name && strcmp(name, "chain_push_barrier") == 0) return g_chain_push.exchange(value);
  if (name && strcmp(name, "head_tail_split") == 0) return g_head_tail_split.exchange(value != 0);
  if (name && strcmp(name, "chain_part_rowmajor") == 0) return g_chain_part_rm.exchange(value != 0);
  if (name && strcmp(name, "head_chain_rows") == 0) return g_head_chain_rows.exchange(value < 2 ? 2 : value);
  if (name && strcmp(name, "head_fuse_tail") == 0) return g_head_fuse_tail.exchange(value != 0);
  if (name && strcmp(name, "dw_bulk") == 0) return g_dw_bulk.exchange(value != 0);
  if (name && strcmp(name, "head_conv_walk") == 0) return g_head_conv_walk.exchange(value != 0);
  if (name && strcmp(name, "enc_splitk") == 0) return g_enc_splitk.exchange(value != 0);
  if (name && strcmp(name, "chain_krot") == 0) return g_chain_krot.exchange(value < 0 ? 0 : value);
  if (name && strcmp(name, "chain_keep_layers") == 0) return g_chain_keep_layers.exchange(value < 0 ? 0 : value);
  if (name && strcmp(name, "chain_wstages") == 0) return g_chain_wstages.exchange(value < 2 ? 2 : value);
  if (name && strcmp(name, "block_pipe") == 0) return g_block_pipe.exchange(value != 0) ? 1 : 0;
  if (name && strcmp(name, "fused_block") == 0) return g_fused_block.exchange(value != 0) ? 1 : 0;
  vv_set_error("vv_set_option: unknown option '%s'", name ? name : "(null)");
  return VV_ERR_INVALID;
}

int vv_linear(const void* A, int lda, const void* W, int ldw, int M, int N, int K, int mode,
              void* out, int ldo, const float* bias, const float* gamma, const void* resid,
              int ldr, const void* gate, int ldg, int bn, int splits, int cluster, int stages,
              void* stream) {
  GemmArgs a;
  a.A = static_cast<const bf16*>(A); a.lda = lda;
  a.B = static_cast<const bf16*>(W); a.ldb = ldw;
  a.M = M; a.N = N; a.K = K; a.mode = mode;
  a.out = out; a.ldo = ldo; a.bias = bias; a.gamma = gamma;
  a.resid = resid; a.ldr = ldr;
  a.gate = static_cast<const bf16*>(gate); a.ldg = ldg;
  a.bn = bn; a.splits = splits; a.cluster = cluster; a.stages = stages;
  a.swap = 2;
  gemm_plan(a, vv_num_sms());
  return gemm_launch(a, static_cast<cudaStream_t>(stream));
}

int vv_debug_trace(long long* out, int cap) {
  const int sel = g_trace_select.load();
  return sel == 4 ? pipe_debug_trace(out, cap) : sel == 3 ? ffn_debug_trace(out, cap) : sel == 2 ? chain_debug_trace(out, cap) : sel == 1 ? block_debug_trace(out, cap) : gemm_debug_trace(out, cap);
}

int vv_timing_enable(int on) {
  gemm_timing_enable(on != 0);
  return VV_OK;
}

int vv_timing_fetch(VVLaunchRec* out, int cap) {
  static_assert(sizeof(VVLaunchRec) == sizeof(GemmLaunchRec), "record layout");
  return gemm_timing_fetch(reinterpret_cast<GemmLaunchRec*>(out), cap);
}

}  // extern "C"
