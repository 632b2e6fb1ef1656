/* vvb200.h — C ABI of libvvb200.so: VibeVoice's per-frame acoustic generation path
 * (diffusion head + DPM-Solver++ loop + sigma-VAE acoustic decoder + streaming conv cache)
 * as hand-written sm_100a CUDA.
 *
 * The reference (lzhgus/VibeVoice) has no FFI for this path: its boundary is three Python
 * methods.  Each entry point below names the reference interface it replaces
 * (paths relative to the reference repo).  The Python host mirror (vibevoice_b200/) binds
 * these with ctypes; INTEGRATION.md shows the reference-side stub.
 *
 * Conventions
 *   - every function returns 0 on success or a negative VVStatus; vv_last_error() gives
 *     the reason (thread-local).  There is no CPU fallback and nothing throws.
 *   - pointers are raw CUDA device pointers unless marked "host"; the caller owns inputs,
 *     outputs and the stream; the library owns packed weights, schedule tables, scratch
 *     and the cache arena.  Calls are asynchronous on `stream` and never host-synchronise
 *     (except *_create, which upload weights, and vv_head_set_schedule).
 *   - `stream` is a cudaStream_t passed as void*.
 *   - dtype codes: 0 = bf16, 1 = fp32.
 */
#ifndef VVB200_H_
#define VVB200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define VV_API __attribute__((visibility("default")))
#else
#define VV_API
#endif

#define VV_DTYPE_BF16 0
#define VV_DTYPE_F32 1

#define VV_ALGO_DPMPP 0     /* "dpmsolver++"      (vibevoice/schedule/dpm_solver.py:676,757) */
#define VV_ALGO_SDE_DPMPP 1 /* "sde-dpmsolver++"  (vibevoice/schedule/dpm_solver.py:680,785) */

typedef struct VVHead VVHead;
typedef struct VVDec VVDec;
typedef struct VVCache VVCache;
typedef struct VVDec VVEnc;        /* an encoder handle shares the decoder's handle type (and its cache arena type) */
typedef struct VVConn VVConn;

VV_API int vv_version(void);
VV_API const char* vv_last_error(void);

/* ---- generic fused linear (the tcgen05 GEMM every layer uses) ------------------------
 * out = epilogue(A[M,K] @ W[N,K]^T); A, W bf16 row-major.  `mode` is a GemmMode value
 * (vibevoice_b200/csrc/vv_gemm.cuh).  Replaces torch.nn.functional.linear call sites
 * (modular_vibevoice_diffusion_head.py:59-62,111-113; modular_vibevoice_tokenizer.py:588-596).
 * bn / splits / cluster / stages = 0 lets the library plan the launch (tile width, split-K
 * factor, CTAs per cluster multicasting the A tile, depth of the TMA ring).  For mode 4
 * (split-K partials) `out` must hold splits*M*ldo floats. */
VV_API int vv_linear(const void* A, int lda, const void* W, int ldw, int M, int N, int K, int mode,
              void* out, int ldo, const float* bias, const float* gamma, const void* resid,
              int ldr, const void* gate, int ldg, int bn, int splits, int cluster, int stages,
              void* stream);
VV_API int vv_linear_plan(int M, int N, int K, int mode, int* bn, int* splits, int* cluster, int* stages);

/* ---- split-K linear that allocates nothing --------------------------------------------------
 * No counterpart in the reference: this entry point was added for callers
 * that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 *   out[M,N] = A[M,K] @ W[N,K]^T + bias      A, W bf16 row-major; bias fp32 [N] or NULL; out bf16 (0) or fp32 (1)
 * K is split over `splits` CTAs per output tile (0 = choose, at most 32); the fp32 partial sums live in a file-scope
 * device array of VV_SPLITK_STATIC_MAX_FLOATS floats (4 MiB on every GPU that loads the library, held until exit)
 * instead of a cudaMalloc'ed buffer, and a second kernel sums them in a fixed order.  The call makes no cudaMalloc,
 * cudaFree or other allocating call; inputs, output and stream are the caller's.  Limits and caveats:
 *   - splits*M*N must be <= VV_SPLITK_STATIC_MAX_FLOATS, else VV_ERR_INVALID (-1) with the sizes in vv_last_error();
 *     N and K multiples of 8, ldo >= N with 16-byte aligned rows;
 *   - every call on a GPU shares the one array, which keeps the last call's partials: the region a call uses is
 *     cleared on `stream` first, and calls that OVERLAP on the GPU (different streams, no dependency) race on it -
 *     serialise them;
 *   - vv_timing_enable(1) must be off (per-launch timing creates CUDA events). */
#define VV_SPLITK_STATIC_MAX_FLOATS (1 << 20)
VV_API int vv_linear_splitk_static(const void* A, int lda, const void* W, int ldw, int M, int N, int K, int splits,
                                   const float* bias, void* out, int ldo, int out_dtype, void* stream);

/* Library-wide switches (process-wide; tuning and A/B measurement - the defaults are the fast path).  Every name:
 *   "pdl"              1|0    programmatic dependent launch between consecutive kernels of a call              (1)
 *   "head_chain"       0|1    vv_head_sample: 0 one launch per op, 1 (or 2) all solver steps in one persistent
 *                             cooperative launch (up to "head_chain_rows" rows; every head, VibeVoice-Large included)  (1)
 *   "fused_block"      1|0    one-launch Block1D for the C = 32/64/128 stages (else mixer + two GEMMs)          (1)
 *   "block_pipe"       1|0    that launch is the pipelined kernel k_block_pipe (vv_blockp.cu) instead of
 *                             k_block_fused (vv_block.cu); same arithmetic, bit-equal results                   (1)
 *   "stage_chain"      k      the C >= 512 stages at one new time step per utterance run as ONE persistent
 *                             launch when rows >= k; 0: never                                                    (1)
 *   "stage_tiled"      rows   the C >= 512 stages at SEVERAL steps per utterance (n * steps <= rows) run as one
 *                             persistent launch too (mixer tiles | FFN1 | FFN2 + residual per block); 0: never;
 *                             measured slower on B200 (1.388 vs 1.305 ms per decoder step)                      (0)
 *   "dw_rows"          r      time steps per CTA of the mixer kernel of the C >= 256 stages; 0: 8192 / C
 *                             (measured: smaller tiles are slower, 1.31 -> 1.34 / 1.38 / 1.47 / 1.72 ms at 8/4/2/1) (0)
 *   "gemm_persistent"  1|0    persistent kernel (double-buffered TMEM accumulator) for dense GEMM launches      (1)
 *   "gemm_stream_bn32" 1|0    32-wide tiles for weight-streaming GEMMs that would leave < 2 CTAs per SM         (1)
 *   "gemm_swap"        1|0    split-K launches with M <= 64: weights on the MMA's 128-row side (measured: no gain) (0)
 *   "chain_push_barrier" 2|1  cross-proxy fences of the persistent phase engine's grid barrier (k_chain): 2 = limited to global
 *                             memory; 1 = unqualified fence.proxy.async (0.4 us per barrier slower; kept for A/B runs)    (2)
 *   "head_tail_split"  0|1    that row-pair phase shares every pair between TWO CTAs (half of the columns each, three small
 *                             exchanges through global memory) instead of one CTA per pair; measured: no gain (0.912 vs
 *                             0.905 ms - an exchange through global memory costs ~2.5 k clk, a grid barrier ~2 k)         (0)
 *   "chain_part_rowmajor" 1|0 the head program's split-K partials are stored [M][splits][H], so a row phase fetches a row's
 *                             partials with one bulk copy (4 copies per row instead of 11); measured: neutral            (1)
 *   "head_chain_rows"  r      vv_head_sample runs as the persistent kernel up to r rows (2 x utterances); above it launch by
 *                             launch.  Measured at 128 / 192 / 256 utterances: 1.41 / 1.94 / 2.36 ms persistent vs 1.57 /
 *                             2.60 / 2.58 ms per launch                                                                      (512)
 *   "head_fuse_tail"   1|0    the solver-step boundary of the head program as ONE row-pair phase (final layer, CFG, solver
 *                             update, input projection, first norm) or as five phases; measured 0.934 vs 0.986 ms         (1)
 *   "dw_bulk"          1|0    the C >= 256 stages' mixer kernel brings its input tile into shared memory with bulk copies
 *                             first (k_block_dw_bulk) instead of reading rows pass by pass (k_block_dw); bit-equal        (1)
 *   "head_conv_walk"   1|0    the decoder's output conv (C -> 1, k = 7) as the register-blocked kernel (4 lanes per output,
 *                             16 outputs per lane group, weights in registers) instead of the shared-memory staged one  (1)
 *   "enc_splitk"       1|0    split-K + row-wise reduce for the encoder's long-K small-M GEMMs                  (1)
 *   "chain_stages"     k>=2   depth of the persistent kernel's activation ring (32 KB stages; rounded up to even)  (3)
 *   "chain_wstages"    k>=2   cap on the depth of its weight ring (16 KB stages; rounded down to even)        (12)
 *   "chain_krot"       r      work item on N tile j starts at k-block (j*r) mod nkb; 0: all start at 0          (0)
 *   "chain_keep_layers" k     vv_head_sample's first k layers' weights are loaded evict-last (kept in L2 across solver
 *                             steps); measured: no gain (1.099 -> 1.109-1.114 ms)                                (0)
 *   "launch_log"       1|0    diagnostic: bracket every launch with CUDA events; switching it off prints kernel name
 *                             and device time per launch to stderr (serialises the launches)                     (0)
 *   "trace_select"     0..4   which kernel's clock stamps vv_debug_trace returns (trace builds only)            (0)
 * Options that shape persistent programs ("chain_*") apply to programs built afterwards.  Returns the previous value,
 * or a negative status for an unknown name. */
VV_API int vv_set_option(const char* name, int value);

/* ---- diffusion head ------------------------------------------------------------------
 * Replaces VibeVoiceDiffusionHead (modular_vibevoice_diffusion_head.py:191-280) and the
 * sampling loop VibeVoiceForConditionalGenerationInference.sample_speech_tokens
 * (modeling_vibevoice_inference.py:712-725) incl. DPMSolverMultistepScheduler.step
 * (vibevoice/schedule/dpm_solver.py:935-1022). */
typedef struct VVHeadDesc {
  int hidden;       /* H: config.hidden_size (1536 / 3584) */
  int ffn;          /* F: int(H * head_ffn_ratio) */
  int latent;       /* L: config.latent_size (64) */
  int layers;       /* config.head_layers (4) */
  int freq_dim;     /* TimestepEmbedder.frequency_embedding_size (256) */
  float eps;        /* config.rms_norm_eps */
  int weight_dtype; /* dtype of the weight pointers passed to vv_head_create (0 bf16, 1 f32) */
  int max_rows;     /* largest M (= 2n) a call will use; scratch is sized for it */
} VVHeadDesc;

/* weights: device pointers in this order (state-dict names):
 *   0 noisy_images_proj.weight [H,L]      1 cond_proj.weight [H,H]
 *   2 t_embedder.mlp.0.weight [H,freq]    3 t_embedder.mlp.2.weight [H,H]
 *   then per layer i: 4+5i norm.weight [H], +1 adaLN_modulation.1.weight [3H,H],
 *        +2 ffn.gate_proj.weight [F,H], +3 ffn.up_proj.weight [F,H], +4 ffn.down_proj.weight [H,F]
 *   then final_layer.adaLN_modulation.1.weight [2H,H], final_layer.linear.weight [L,H] */
VV_API int vv_head_create(const VVHeadDesc* desc, const void* const* weights, int n_weights,
                   VVHead** out);
VV_API void vv_head_destroy(VVHead* h);

/* Replaces set_timesteps (dpm_solver.py:321-423) + the per-step scalar math of
 * dpm_solver.py:669-677,738-793.  t_values (host, S): timesteps already rounded to the
 * model dtype (modeling_vibevoice_inference.py:720); sigmas (host, S+1) as produced by
 * set_timesteps (last = 0). */
VV_API int vv_head_set_schedule(VVHead* h, int S, const float* t_values, const float* sigmas, int algo,
                         void* stream);

/* Drop-in for prediction_head.forward(noisy_images[M,L], timesteps[M], condition[M,H])
 * (modular_vibevoice_diffusion_head.py:254-280).  io_dtype applies to z, cond and out;
 * t is fp32 [M] (device). */
VV_API int vv_head_forward(VVHead* h, const void* z, const float* t, const void* cond, void* out, int M,
                    int io_dtype, void* stream);

/* Drop-in for sample_speech_tokens(cond[n,H], neg[n,H], cfg) -> [n,L].
 * noise: fp32 [2n,L], the reference's torch.randn(2n, L) draw (rows n..2n are unused, as in
 * the reference).  sde_noise: fp32 [S,n,L] or NULL (required for VV_ALGO_SDE_DPMPP). */
VV_API int vv_head_sample(VVHead* h, const void* cond, const void* neg, const float* noise,
                   const float* sde_noise, float cfg, void* out_latent, int n, int io_dtype,
                   void* stream);

/* ---- acoustic decoder + streaming cache ----------------------------------------------
 * Replaces VibeVoiceAcousticTokenizerModel.decode / TokenizerDecoder
 * (modular_vibevoice_tokenizer.py:816-951,1099-1108) and VibeVoiceTokenizerStreamingCache
 * (modular_vibevoice_tokenizer.py:193-256). */
#define VV_DEC_MAX_STAGES 8
typedef struct VVDecDesc {
  int vae_dim;                      /* 64 */
  int n_stages;                     /* len(depths) = 7 */
  int channels[VV_DEC_MAX_STAGES];  /* per-stage width: n_filters * 2^(n_stages-1-i) */
  int depths[VV_DEC_MAX_STAGES];    /* Block1D count per stage */
  int ratios[VV_DEC_MAX_STAGES];    /* upsample stride into stage i (ratios[0] unused = 1) */
  int kernel;                       /* 7: stem / depthwise / head conv kernel size */
  float eps;                        /* layernorm_eps */
  int weight_dtype;
  int max_rows;                     /* largest n * T (latent frames) per call */
} VVDecDesc;

/* weights: device pointers, order:
 *   stem conv weight [C0,vae,k], bias [C0];
 *   for stage i: (i>0: convtr weight [C(i-1),C(i),2r], bias [C(i)]), then per block:
 *      norm.weight[C], mixer dw weight [C,1,k], mixer bias [C], gamma [C],
 *      ffn_norm.weight [C], ffn.linear1.weight [4C,C], bias [4C], ffn.linear2.weight [C,4C],
 *      bias [C], ffn_gamma [C];
 *   head conv weight [1,C_last,k], bias [1]. */
VV_API int vv_dec_create(const VVDecDesc* desc, const void* const* weights, int n_weights, VVDec** out);
VV_API void vv_dec_destroy(VVDec* d);

/* Offline decode (use_cache=False; zero history).  latents: [n,T,vae] when time_major != 0,
 * else [n,vae,T].  wave: [n,1,hop*T].  Applies z/scale - bias first
 * (modeling_vibevoice_inference.py:651); pass scale=1, bias=0 for a plain decode. */
VV_API int vv_dec_decode(VVDec* d, const void* latents, int n, int T, int time_major, float scale,
                  float bias, void* wave, int io_dtype, void* stream);

/* Streaming cache arena: max_slots utterances, zero-initialised (== no history). */
VV_API int vv_cache_create(VVDec* d, int max_slots, VVCache** out);
VV_API void vv_cache_destroy(VVCache* c);
/* set_to_zero(sample_indices) (modular_vibevoice_tokenizer.py:234-241); slot_idx: device int64 [k] */
VV_API int vv_cache_zero(VVCache* c, const int64_t* slot_idx, int k, void* stream);
VV_API int vv_cache_zero_all(VVCache* c, void* stream);
/* Streaming decode step (use_cache=True): row i uses cache slot slot_idx[i] (device int64 [n]). */
VV_API int vv_dec_step(VVDec* d, VVCache* c, const void* latents, const int64_t* slot_idx, int n, int T,
                int time_major, float scale, float bias, void* wave, int io_dtype, void* stream);

/* ---- semantic tokenizer encoder step ---------------------
 * Replaces VibeVoiceSemanticTokenizerModel.encode(...).mean / TokenizerEncoder
 * (modular_vibevoice_tokenizer.py:687-813,1118-1170), which generate() calls once per frame on the
 * chunk the decoder just produced (modeling_vibevoice_inference.py:673-679).
 * desc: vae_dim = output dim (128); channels[i] = n_filters * 2^i; ratios[i] = stride of the conv INTO
 * stage i (ratios[0] unused); max_rows = largest n * T (frames of hop samples) per call.
 * weights: device pointers in TokenizerEncoder state-dict order:
 *   stem conv weight [C0,1,k], bias [C0];
 *   for stage i: (i>0: strided conv weight [C(i),C(i-1),2r], bias [C(i)]), then the ten Block1D tensors
 *   per block as for vv_dec_create;  head conv weight [dim,C_last,k], bias [dim].
 * The cache is a vv_cache_create(enc, ...) arena; vv_cache_zero / vv_cache_zero_all apply. */
VV_API int vv_enc_create(const VVDecDesc* desc, const void* const* weights, int n_weights, VVEnc** out);
VV_API void vv_enc_destroy(VVEnc* e);
/* Offline encode (use_cache=False; zero history): audio [n, 1, hop*T] -> out [n, T, dim]; chunked in time
 * internally (max_rows >= n). */
VV_API int vv_enc_encode(VVEnc* e, const void* audio, int n, int T, void* out, int io_dtype, void* stream);
/* Voice-prompt prefill (acoustic_tokenizer.encode without a cache,
 * modeling_vibevoice_inference.py:154, on TokenizerEncoder.forward :774-813): audio [n, n_samples] of ANY length ->
 * out [n, ceil(n_samples/hop), dim].  The last, partial frame is completed with zeros stage by stage exactly as the
 * non-streaming SConv1d's "extra padding" does (modular_vibevoice_tokenizer.py:127-133, :393-401); the call is
 * chunked in time internally, so n_samples is bounded only by max_rows >= n. */
VV_API int vv_enc_prefill(VVEnc* e, const void* audio, int n, long long n_samples, void* out, int io_dtype, void* stream);
/* Streaming step (use_cache=True): row i uses cache slot slot_idx[i] (device int64 [n]). */
VV_API int vv_enc_step(VVEnc* e, VVCache* c, const void* audio, const int64_t* slot_idx, int n, int T, void* out,
                int io_dtype, void* stream);

/* ---- SpeechConnector -------------------------------------
 * Replaces SpeechConnector.forward (modeling_vibevoice.py:58-69): fc1 -> LlamaRMSNorm(eps) -> fc2.
 * weights: fc1.weight [out,in], fc1.bias [out], norm.weight [out], fc2.weight [out,out], fc2.bias [out].
 * vv_conn_forward: x [M, in_dim] (io_dtype) -> out [M, out_dim] bf16; when addend (bf16 [M, out_dim]) is
 * given it is added in the last epilogue (acoustic_embed + semantic_embed, modeling_vibevoice_inference.py:684). */
VV_API int vv_conn_create(int in_dim, int out_dim, float eps, int max_rows, int weight_dtype,
                   const void* const* weights, int n_weights, VVConn** out);
VV_API void vv_conn_destroy(VVConn* c);
VV_API int vv_conn_forward(VVConn* c, const void* x, int M, int io_dtype, const void* addend, void* out, void* stream);

/* Prompt latents -> acoustic features: VibeVoiceTokenizerEncoderOutput.sample (modular_vibevoice_tokenizer.py:1018-1046)
 * fused with (latents + speech_bias_factor) * speech_scaling_factor (modeling_vibevoice_inference.py:156-159, :167-170).
 * mean, noise, out: [n, T, dim] in io_dtype; the caller draws the noise (randn_like(mean), and for 'gaussian'
 * row_noise = randn(n)) so that its generator state advances as in the reference.
 *   dist 'fix':      noise given, row_noise NULL, std = fix_std
 *   dist 'gaussian': noise and row_noise given, std = fix_std / 0.8; std_out (optional, [n]) receives row_noise * std
 *   no sampling:     noise NULL (out = (mean + bias) * scale)
 * Each elementwise op rounds to io_dtype, as the eager reference does. */
VV_API int vv_latent_sample(const void* mean, const void* noise, const void* row_noise, float std, float bias, float scale,
                     int n, int T, int dim, int io_dtype, void* out, void* std_out, void* stream);

/* ---- Frame-loop bookkeeping on the device ------------------
 * Replaces the host-side row selection of generate() (modeling_vibevoice_inference.py:588 `torch.arange(batch)[~finished
 * & (next_tokens == speech_diffusion_id)]`, a sync per frame), the per-row `.item()` + list append of the audio chunks
 * (:653-658) and the indexed embedding update (:687).  The whole batch runs every frame at a fixed launch shape;
 * inactive rows are pointed at a spare cache slot and dropped by the masked writes.
 *
 * vv_frame_select: active[i] = !finished[i] && next_tokens[i] == token_id   (either input may be NULL: no finished rows /
 *   active = !finished).  Outputs, each optional: slots_out[i] = active ? i : trash_slot (int64, the slot_idx argument of
 *   vv_dec_step / vv_enc_step / vv_cache_zero); active_out[i] (1 byte, a torch.bool tensor); compact_out = the active row
 *   numbers in ascending order followed by -1 (the reference's diffusion_indices); count_out = how many.  `finished` is
 *   1 byte per row (torch.bool).
 * vv_ring_append: for every active row, ring[row, pos[row]] = chunk[row] (one frame of `hop` samples, io_dtype) and
 *   pos[row] += 1.  ring is [n, cap_frames, hop]; a row that is full drops the frame but keeps counting, so
 *   pos[row] > cap_frames tells the caller afterwards.
 * vv_embed_scatter: dst[row, :] = emb[row, :] for active rows (emb bf16 [n, H]; dst [n, H] bf16 or fp32). */
VV_API int vv_frame_select(const int64_t* next_tokens, const void* finished, int n, int64_t token_id, int64_t trash_slot,
                    int64_t* slots_out, void* active_out, int64_t* compact_out, int* count_out, void* stream);
VV_API int vv_ring_append(const void* chunk, const void* active, int n, int hop, int io_dtype, void* ring, int* pos,
                   int cap_frames, void* stream);
VV_API int vv_embed_scatter(const void* emb, const void* active, int n, int H, void* dst, int dst_dtype, void* stream);

/* Measurement aid (bench.py's live roofline): while enabled, every GEMM launch and every launch of
 * the persistent solver-loop kernel is bracketed by CUDA events on its own stream.
 * vv_timing_fetch synchronises on them, writes up to `cap` records and clears the list; returns the
 * number written.  kind 0: one GEMM (M, N, K, bn, splits, mode; bytes = 0, the caller derives them);
 * kind 1: one persistent head launch (M = rows, N = phases, K = GEMM phases; bytes = compulsory HBM
 * bytes of the whole launch: every weight and adaLN-table row once per phase that uses it; activations
 * and split-K partials are L2-resident scratch and not counted). */
typedef struct VVLaunchRec { int M, N, K, bn, splits, mode; float ms; int kind; long long bytes; } VVLaunchRec;
VV_API int vv_timing_enable(int on);
VV_API int vv_timing_fetch(VVLaunchRec* out, int cap);

/* Debug builds (make TRACE=1) only: per-CTA clock64 stamps of the last GEMM launch, [160][48] int64.
 * Returns the number of elements written; 0 in a normal build. */
VV_API int vv_debug_trace(long long* out, int cap);

/* Measurement probes (tcgen05.commit cost, TMA streaming rate) are not part of the product library: they are declared in
 * include/vvb200_probe.h and compiled only with `make -C vibevoice_b200/csrc PROBE=1`. */
/* Number of kernels the library has launched since load (bench.py's gpu_launches). */
VV_API long long vv_launch_count(void);

/* Number of host synchronisations the LIBRARY itself has made since load: building a persistent program for a batch
 * size it has not seen (vv_head_sample, vv_dec_step, vv_enc_step: once per size), vv_head_set_schedule, dropping the
 * program cache, *_create.  A steady-state frame loop must not move it (tests/test_dropin.py).  No counterpart in the
 * reference, whose loop synchronises dozens of times per frame (.tolist() / .item()). */
VV_API long long vv_host_sync_count(void);

#ifdef __cplusplus
}
#endif
#endif /* VVB200_H_ */
