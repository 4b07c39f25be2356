/* echo_b200.h -- C ABI of libecho_b200.so (hand-written sm_100a kernels for the Echo-TTS
 * sampling hot path).  Plain pointers and sizes only; every pointer is a DEVICE pointer owned
 * by the caller (PyTorch tensors on the Python side), every call is asynchronous on `stream`,
 * nothing is allocated inside.  Return value: 0 ok, <0 error (echo_last_error() has the text).
 *
 * The upstream project (CoreBedtime/echo-tts) is pure Python/PyTorch and has no FFI; each entry
 * point below names the reference code it replaces (file:line in the upstream repo) -- the
 * binding a maintainer would add is the ctypes stub shown in INTEGRATION.md.
 */
#ifndef ECHO_B200_H
#define ECHO_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* echo_stream_t; /* cudaStream_t */

const char* echo_last_error(void);
int echo_version(void);
unsigned long long echo_launch_count(void); /* kernels launched by this library so far */
void echo_reset_launch_count(void);

/* ------------------------------------------------------------------------------------------
 * bf16 tensor-core GEMM with implicit-conv taps and fused epilogues.
 *   acc[b, m, n] = sum_t sum_k A[b, m + tap_shift[t], k] * B[n, t*k_tap + k]     (fp32 accumulate)
 * Rows outside [-a_ctx, rows) read as zero (causal left padding of CausalConvNet,
 * autoencoder.py:285-289).  ONE kernel: the persistent CTA-pair tcgen05.mma kernel with TMEM accumulators
 * and TMA loads (csrc/gemm2.cu); `impl` is kept in the descriptor for ABI stability (0 or 2: that kernel).
 * Replaces every nn.Linear / Conv1d / ConvTranspose1d on the path: model.py:217-224, 266, 308;
 * autoencoder.py:285-316, 365-367, 667, 704, 717.
 * ------------------------------------------------------------------------------------------ */
enum { ECHO_EPI_GENERIC = 0, ECHO_EPI_SECTIONS = 1, ECHO_EPI_SWIGLU = 2 };
enum { ECHO_ACT_NONE = 0, ECHO_ACT_GELU = 1, ECHO_ACT_SILU = 2, ECHO_ACT_SIGMOID = 3, ECHO_ACT_TANH = 4 };
enum { ECHO_SEC_PLAIN = 0, ECHO_SEC_NORM_ROPE = 1, ECHO_SEC_SIGMOID = 2, ECHO_SEC_ROPE = 3 };

typedef struct echo_gemm_section {
  int32_t type;             /* ECHO_SEC_* */
  int32_t _pad;
  void* out;                /* bf16 [batch*rows, ld_sec_out] */
  const float* norm_weight; /* fp32 [section_width] (per-head RMSNorm weight, model.py:192-193) */
} echo_gemm_section;

typedef struct echo_gemm_desc {
  const void* a;            /* bf16 [batch][rows][lda] */
  const void* b;            /* bf16 [n][ldb], ldb >= taps*k_tap */
  int64_t lda, a_batch_stride, ldb;
  int32_t rows, batch, n, k_tap, taps;
  int32_t tap_shift[8];
  int32_t epilogue;         /* ECHO_EPI_* */
  int32_t impl;             /* 0 or 2: the CTA-pair tcgen05 kernel (the only GEMM kernel); 3: the same, and the library may
                             * schedule ragged launches stream-K through `workspace`; 4: the same with the plain tile
                             * schedule (no half-width last round; results are bit-identical to 2, kept for A/B and
                             * the tests); 5: the same with the transposing GENERIC epilogue where 2 would take the row-per-thread
                             * TMA epilogue (bias [+ residual] -> fp32 / one bf16 output: bit-identical, for A/B and the
                             * tests); anything else is refused */
  int32_t block_n;          /* 0 = auto; 64 / 128 / 256 */
  int32_t act;              /* GENERIC: activation applied after bias */
  /* GENERIC: v = act(acc + bias[n]) * gate[(m / rows_per_gate) * gate_ld + n] + resid[m, n]
   *          out_f32 / out_bf16 <- v ; out2_bf16 <- snake(v, snake_alpha[n])  (any may be NULL) */
  const float* bias;
  const float* gate;
  int64_t gate_ld;
  int32_t rows_per_gate;
  int32_t k_splits;         /* impl 2, GENERIC only: > 1 splits the K loop over k_splits work units per tile; unit s writes
                             * its partial result (bias only in unit 0; act must be NONE, resid NULL) to
                             * out_f32 + s * split_stride -- the consumer (echo_rmsnorm_mod / echo_cfg_euler_tail `delta`
                             * arguments) sums them in a fixed order, so the result is deterministic.  0 or 1 = off. */
  const float* resid;
  int64_t ld_resid;
  float* out_f32;
  int64_t ld_out_f32;
  void* out_bf16;
  int64_t ld_out_bf16;
  void* out2_bf16;
  int64_t ld_out2_bf16;
  const float* snake_alpha;
  /* SECTIONS: n = n_sections * section_width; per-head RMSNorm (+ interleaved-pair RoPE on heads
   * < rope_heads at position pos_offset + (m % rows_per_seq) * pos_stride) / plain / sigmoid.
   * Replaces model.py:221-232, 199-202, 17-24, 264 (sigmoid of the gate) and 270-293. */
  int32_t n_sections, section_width, head_dim, rope_heads;
  int32_t pos_offset, pos_stride, rows_per_seq;
  int32_t a_ctx;            /* streaming (S1-DAC decode stream): rows of left context that sit in memory BEFORE row 0 of every
                             * batch entry of `a`; rows -a_ctx..-1 are read instead of zero-filled (the causal padding of
                             * the first call is an all-zero context).  a_batch_stride must span them.  0 = none. */
  float eps;
  int32_t b_static;         /* 1 = `b` is a load-time tensor (weights): nothing launched on `stream` since the last
                             * synchronisation or non-library launch writes it.  The kernel then requests its first B
                             * tiles while the previous launch is still draining (programmatic dependent launch) instead of
                             * after it; results are identical.  0 = no such promise (always safe). */
  const float* rope_cos;    /* fp32 [max_pos][head_dim/2] */
  const float* rope_sin;
  int64_t ld_sec_out;
  echo_gemm_section sections[4];
  int64_t split_stride;     /* elements between the k_splits partial outputs */
  /* SWIGLU: B rows interleaved in groups of 64 (w1 | w3); out_bf16[m, j] = silu(a)*b (model.py:308) */
  void* workspace;          /* optional: device scratch of >= echo_gemm_workspace_bytes(), zeroed once by the
                             * caller, not shared between concurrently running streams.  With impl 3 the library may
                             * schedule the launch stream-K (equal k-block ranges per CTA pair, partial accumulators
                             * exchanged through this buffer in a fixed order: deterministic, same epilogues). */
  int64_t workspace_bytes;
} echo_gemm_desc;

int echo_gemm_bf16(const echo_gemm_desc* d, echo_stream_t stream);
long long echo_gemm_workspace_bytes(void);

/* NO COUNTERPART IN THE REFERENCE.  Allocation-free split-K GEMM for
 * callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it):
 * echo_gemm_bf16 in its split-K form (d->k_splits >= 2, GENERIC epilogue, no activation, fp32 output) with the partial
 * products [k_splits][rows * batch][n] fp32 held in a file-scope `static __device__` array of
 * ECHO_NOALLOC_SCRATCH_BYTES instead of a caller-lent buffer, followed by a fixed-order reduction:
 *     out_f32[m, n] = (resid ? resid[m, n] : 0) + sum_s gate * (A_s W_s^T + bias [s == 0])
 * i.e. the result echo_gemm_bf16 gives without k_splits (up to fp32 summation order).  d->split_stride / d->workspace
 * are ignored; resid may alias out_f32.  Calls no cudaMalloc / cudaFree / allocating library routine; every GPU that
 * loads the library holds the array until exit.  Requests with k_splits * rows * batch * n * 4 > ECHO_NOALLOC_SCRATCH_BYTES
 * are refused (ECHO_ERR_INVALID, message in echo_last_error()).  All calls on a device share the one array, which keeps
 * the previous call's contents (nothing needs clearing: a call overwrites every element it reads); calls that overlap on
 * the GPU race -- serialise them (one stream, or graph edges between them). */
#define ECHO_NOALLOC_SCRATCH_BYTES (32LL << 20)
int echo_gemm_bf16_splitk_noalloc(const echo_gemm_desc* d, echo_stream_t stream);
long long echo_noalloc_scratch_bytes(void);

/* Debug aid (no reference counterpart; DEBUG BUILDS ONLY: `make -C echo_tts_b200/csrc DEBUG=1` compiles it in with
 * -DECHO_DEBUG, the product library returns ECHO_ERR_UNSUPPORTED for a non-NULL buffer and contains none of it):
 * while `device_buf` (>= 32 int64 per CTA of the launch) is non-NULL, every GEMM launch records clock64() per CTA: [0] start, [1] setup done, [2] first operands landed, [6] roles
 * done, [7] exit, and for its first six tiles t: [8+4t] accumulator stage free, [9+4t] MMAs issued, [10+4t]
 * accumulator ready, [11+4t] epilogue done.  Pass NULL to switch it off.  While tracing, the environment variables
 * ECHO_GEMM_NOLOAD=1 (no TMA after the first ring fill) and ECHO_GEMM_SKIP=A|B (that operand is never loaded) turn the
 * launch into a timing experiment (results are garbage): they separate the MMA rate from each operand's feed. */
int echo_debug_gemm_trace(long long* device_buf);

/* ------------------------------------------------------------------------------------------
 * fp32 CUDA-core GEMM for the accuracy-critical small projections:
 *   out[z, m, n] = post( sum_k pre(A[z, m, k]) * W[z, n, k] + bias[z, n] + addend[m, n] )
 * Replaces cond_module + LowRankAdaLN (model.py:70-74, 532-538), in_proj (model.py:586),
 * SpeakerEncoder.in_proj (model.py:461-462) and the PCA inverse (inference.py:234).
 * ------------------------------------------------------------------------------------------ */
enum { ECHO_POST_NONE = 0, ECHO_POST_PLUS1 = 1, ECHO_POST_TANH = 2, ECHO_POST_SILU = 3 };
typedef struct echo_sgemm_desc {
  const float* a; int64_t lda, a_zstride;
  const float* w; int64_t ldw, w_zstride;
  const float* bias; int64_t bias_zstride;
  const float* addend; int64_t ld_addend, addend_zstride;
  float* out; int64_t ldo, out_zstride;
  void* out_bf16; int64_t ldo_bf16, out_bf16_zstride;
  int32_t m, n, k, nz;
  int32_t pre_silu, post;
  float out_scale;   /* applied after bias, before addend/post (1/6 of model.py:462; 1/latent_scale) */
  float _pad;
} echo_sgemm_desc;
int echo_sgemm_f32(const echo_sgemm_desc* d, echo_stream_t stream);

/* timestep embedding cat[cos(t w), sin(t w)] (model.py:27-43): t fp32 [n] -> out fp32 [n, size] */
int echo_timestep_embedding(const float* t, int n, int size, float* out, echo_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Row normalisation -> bf16 GEMM operand.
 *   y[m, :] = x[m, :] * rsqrt(mean(x^2) + eps) * scale1[g, :] + shift[g, :],  g = mod_base + m / rows_per_mod
 * (LowRankAdaLN.forward model.py:76-79 with scale1 = 1 + scale; RMSNorm model.py:99-104 with
 * shift = NULL, scale1 = weight; autoencoder.py:726-731).  `step` (device int, nullable) adds
 * step[0] * mod_step_stride rows to the table offset so one CUDA graph serves every Euler step.
 * ------------------------------------------------------------------------------------------ */
int echo_rmsnorm_mod(const float* x, int64_t ldx, int m, int d, float eps, const float* scale1,
                     const float* shift, int64_t mod_ld, int rows_per_mod, const int32_t* step,
                     int64_t mod_step_stride, void* y_bf16, int64_t ldy, echo_stream_t stream);
/* Same, preceded by the deferred residual update x[m, :] += sum_s delta[s * delta_stride + m * ldx + :] (the split-K
 * partial outputs of the previous projection, already multiplied by the AdaLN gate: model.py:385, 388); x is
 * updated in place.  n_delta = 0 is echo_rmsnorm_mod. */
int echo_add_rmsnorm_mod(float* x, int64_t ldx, int m, int d, float eps, const float* delta, int n_delta,
                         int64_t delta_stride, const float* scale1, const float* shift, int64_t mod_ld,
                         int rows_per_mod, void* y_bf16, int64_t ldy, echo_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Flash-style attention over up to four KV segments walked in place (no concat):
 * replaces torch.cat + F.scaled_dot_product_attention + sigmoid gating at model.py:246-264 and
 * 148-157, _concat_kv_caches / _multiply_kv_cache (inference.py:409-428) and the dense window
 * mask of autoencoder.py:762-797.
 * ------------------------------------------------------------------------------------------ */
typedef struct echo_attn_segment {
  const void* k;            /* bf16; key j of kv-row b, head h at k + b*batch_stride + j*row_stride + h*head_dim */
  const void* v;
  int64_t row_stride, batch_stride;
  int32_t len;              /* keys in this segment */
  int32_t kv_batch_mod;     /* kv-row = row % kv_batch_mod (0: kv-row = row): CFG rows share one cache */
  const uint8_t* mask;      /* [rows, mask_stride] key-valid bytes, NULL = all valid */
  int64_t mask_stride;
  const int32_t* extent;    /* per-row 1 + last valid key (tile skipping), NULL = len */
  const float* kv_scale;    /* device scalar multiplying K and V (speaker_kv_scale), NULL = 1 */
  int32_t causal;           /* key j visible to query i iff j <= i + q_offset (+ window: j > i + q_offset - window) */
  int32_t window;
  int32_t q_offset;         /* causal segments: key index of query 0 (keys cached from earlier calls precede it:
                             * incremental encoders, S1-DAC decode stream).  0 = queries and keys aligned. */
  int32_t _pad;
} echo_attn_segment;

typedef struct echo_attn_desc {
  const void* q;            /* bf16 [rows, seq, heads, head_dim] */
  int64_t q_row_stride, q_batch_stride; /* elements between consecutive queries / rows */
  const void* gate;         /* bf16 sigmoid(gate) [rows*seq, heads*head_dim] or NULL */
  void* out;                /* bf16 [rows*seq, heads*head_dim] */
  int32_t rows, seq, heads, head_dim;
  int32_t n_seg;
  int32_t impl;             /* 0 = tcgen05 kernel when eligible (head_dim 128, no sliding window; causal segments with
                             * q_offset >= 0 included), else the mma.sync kernel; 1 = force mma.sync */
  float softmax_scale;
  int32_t kv_splits;        /* tcgen05 kernel: 0 = automatic, 1 = never, 2..4 = that many CTAs per (row, head, query tile),
                             * each a contiguous share of the key tiles (split-KV; needs `workspace`) */
  echo_attn_segment seg[4];
  void* workspace;          /* optional split-KV exchange buffer of >= echo_attention_workspace_bytes(...), zeroed ONCE by
                             * the caller, not shared between streams that run attention concurrently.  With few
                             * (row, head, query tile) groups -- the un-guided 640-token step has 80 for 148 SMs, a
                             * 160-latent streaming block 32 -- the parts' un-normalised (O, m, l) go through it and the
                             * last part to arrive merges them in part order: deterministic. NULL = never split. */
  int64_t workspace_bytes;
} echo_attn_desc;
int echo_attention(const echo_attn_desc* d, echo_stream_t stream);
long long echo_attention_workspace_bytes(int rows, int seq, int heads, int kv_splits);
/* Debug aid (no reference counterpart; debug builds only, as echo_debug_gemm_trace): while non-NULL, every tcgen05
 * attention launch records 64 clock64() stamps per CTA (layout documented at the definition in csrc/attention2.cu). */
int echo_debug_attn_trace(long long* device_buf);

/* extent[r] = 1 + index of last non-zero byte of mask[r, :len] (0 if none) */
int echo_mask_extent(const uint8_t* mask, int rows, int len, int64_t stride, int32_t* extent,
                     echo_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Fused sampler tail: out_norm -> out_proj -> 3-branch CFG combine -> temporal score rescale ->
 * Euler update, fp32, x_t updated in place.  Replaces model.py:601-604 and inference.py:514-530,
 * 431-443, 558.  h: final residual stream fp32 [branches*B*S, d] (branch-major: cond | text-
 * dropped | speaker-dropped).  sched: fp32 [steps][4] = {t, t_next - t, unused, unused} read at
 * row step[0] (device) so the call is CUDA-graph replayable.  v_out (nullable): per-branch
 * velocities fp32 [branches*B*S, latent] for parity tests.
 * ------------------------------------------------------------------------------------------ */
typedef struct echo_tail_desc {
  const float* h; int64_t ldh;
  const float* norm_w; const float* w_out; const float* b_out;  /* [d], [latent, d], [latent] */
  float* x_t;               /* fp32 [B*S, latent] in/out */
  float* v_out;
  float* v_comb_out;        /* nullable: combined (post-rescale) velocity fp32 [B*S, latent] */
  const float* sched; const int32_t* step;
  int32_t positions;        /* B*S */
  int32_t d, latent, branches;
  float eps, cfg_text, cfg_speaker, rescale_k, rescale_sigma;
  int32_t use_rescale;
  int32_t update_x;         /* 0: only write velocities (model.__call__ compatibility path) */
  int32_t n_delta;          /* deferred residual update of the last projection: h += sum_s delta[s * delta_stride + ...] */
  const float* delta;       /* fp32 [n_delta][rows][ldh] or NULL */
  int64_t delta_stride;
  /* Multi-GPU CFG-branch split with the exchange fused into the kernels (peer memory over NVLink, no NCCL call):
   * echo_cfg_euler_tail with branches = 1 additionally stores its velocities at v_peer[0..1] (the same slot of the two
   * other ranks' gather buffers, P2P-mapped) and then adds 1 per CTA to flag_peer[0..1] (system-scope release);
   * echo_cfg_combine_euler first waits until wait_flag[0..1] (this rank's counters for the two other senders) reach
   * wait_value.  All NULL / 0 = off. */
  float* v_peer[2];
  int32_t* flag_peer[2];
  const int32_t* wait_flag[2];
  int32_t wait_value, _pad_tail;
  /* nullable device fp32 [4] = {cfg_text, cfg_speaker, rescale_k, rescale_sigma}: when given it overrides the four
   * immediates above, so one captured CUDA graph serves every guidance setting (use_rescale stays a launch-time flag) */
  const float* guidance;
} echo_tail_desc;
int echo_cfg_euler_tail(const echo_tail_desc* d, echo_stream_t stream);
/* Combine-only variant for the multi-GPU CFG-branch split: d->v_out is INPUT, the three branch
 * velocities fp32 [3][positions][latent] gathered from the branch ranks; h / norm_w / w_out / b_out are ignored.
 * Same combine, rescale and Euler arithmetic as echo_cfg_euler_tail (inference.py:526-530, 431-443, 558). */
int echo_cfg_combine_euler(const echo_tail_desc* d, echo_stream_t stream);
/* CTAs one echo_cfg_euler_tail launch uses for `positions` rows = increments per launch of each flag_peer counter. */
int echo_cfg_tail_ctas(int positions);
int echo_step_advance(int32_t* step, echo_stream_t stream); /* step[0] += 1 */

/* ------------------------------------------------------------------------------------------
 * Whole-step entry points: one velocity evaluation of the DiT (EchoDiT.forward, model.py:563-604: in_proj, n_layers x
 * [AdaLN norm -> q|k|v|gate projection + head RMSNorm + RoPE -> joint attention over self / latent-prefix / text /
 * speaker keys -> gated wo residual -> AdaLN norm -> SwiGLU -> gated w2 residual]) followed, when `tail` is given, by
 * the fused out_norm + out_proj + CFG combine + rescale + Euler update (inference.py:514-558); and the whole Euler
 * loop (inference.py:508-558).  Pure launch sequences over the entry points above (about 7 launches per layer), on
 * `stream`, CUDA-graph capturable; `layers`, `k`/`v` pointer arrays and `has_cfg` are HOST arrays read at call time,
 * everything else is device memory lent by the caller.
 * ------------------------------------------------------------------------------------------ */
typedef struct echo_dit_layer {
  const void* w_qkvg;       /* bf16 [4*d, d]: wq | wk | wv | gate rows (model.py:215-224) */
  const void* w_o;          /* bf16 [d, d] (model.py:266) */
  const void* w13;          /* bf16 [2*ffn, d], w1 / w3 rows interleaved in groups of 64 (model.py:296-308) */
  const void* w2;           /* bf16 [d, ffn] */
  const float* q_norm;      /* fp32 [heads*head_dim] per-head RMSNorm weights (model.py:192-193) */
  const float* k_norm;
} echo_dit_layer;

typedef struct echo_dit_kv { /* cached keys / values of one conditioning modality, all layers */
  const void* const* k;     /* HOST array [n_layers] of device pointers, bf16 [kv_rows, >= len, heads, head_dim] */
  const void* const* v;
  int64_t row_stride, batch_stride;
  int32_t len;              /* keys visible to this forward (0 = none: a latent prefix that does not exist yet) */
  int32_t kv_batch_mod;     /* as echo_attn_segment */
  const uint8_t* mask; int64_t mask_stride;
  const int32_t* extent;
  const float* kv_scale;    /* device fp32 [n_layers] per-layer multiplier (speaker_kv_scale, inference.py:420-428) or NULL */
} echo_dit_kv;

typedef struct echo_dit_step_desc {
  int32_t n_layers, d, heads, head_dim, ffn, latent;
  float eps;
  int32_t rope_len;         /* positions in rope_cos / rope_sin */
  const echo_dit_layer* layers; /* HOST array [n_layers] */
  const float* in_w;        /* fp32 [d, latent] */
  const float* in_b;        /* fp32 [d] */
  const float* rope_cos; const float* rope_sin; /* fp32 [rope_len, head_dim/2] */
  const float* x_in;        /* fp32 [x_rows, seq, latent]; x_rows == rows, or rows % x_rows == 0: CFG branches share x_t */
  int32_t x_rows, rows, seq;
  int32_t pos_offset;       /* RoPE position of the first latent (start_pos, model.py:573-579) */
  const float* mod;         /* fp32 [mod_rows][n_layers][2][3][d] = {shift, 1 + scale, tanh-ed gate} (model.py:60-83) */
  int64_t mod_row_stride;
  int32_t mod_rows;         /* 1 (one t for all rows) or rows */
  int32_t k_splits;         /* > 1: wo / w2 run split-K into `delta`, summed by the next norm / the tail */
  echo_dit_kv prefix, text, speaker;
  float* x;                 /* workspace: residual stream fp32 [rows*seq, d] (result when tail == NULL) */
  void* xn;                 /* bf16 [rows*seq, d] */
  void* qkvg;               /* bf16 [4][rows*seq, d] */
  void* ao;                 /* bf16 [rows*seq, d] */
  void* hmid;               /* bf16 [rows*seq, ffn] */
  float* delta;             /* fp32 [k_splits][rows*seq, d] when k_splits > 1 */
  void* gemm_workspace; int64_t gemm_workspace_bytes; /* as echo_gemm_desc.workspace */
  int32_t gemm_impl, attn_impl; /* echo_gemm_desc.impl / echo_attn_desc.impl for every launch (2 / 0 = product path) */
  const echo_tail_desc* tail;   /* nullable; h, ldh, delta, n_delta, delta_stride are filled in by the library */
  int32_t* pending_out;     /* HOST, nullable: set to 1 when tail == NULL and `delta` still has to be added to x */
  void* attn_workspace; int64_t attn_workspace_bytes; /* as echo_attn_desc.workspace (split-KV); NULL = never split */
} echo_dit_step_desc;
int echo_dit_step(const echo_dit_step_desc* d, echo_stream_t stream);

/* The Euler loop (inference.py:508-558): step i runs *cfg_step (three-branch rows) when has_cfg[i] else *plain_step,
 * with mod, speaker.kv_scale (when non-NULL) and tail->sched advanced by i * the given element strides. */
int echo_dit_sample(const echo_dit_step_desc* cfg_step, const echo_dit_step_desc* plain_step, int n_steps,
                    const uint8_t* has_cfg, int64_t mod_step_stride, int64_t kv_scale_step_stride,
                    int64_t sched_step_stride, echo_stream_t stream);

/* text embedding lookup (model.py:420): ids int32 [n] -> out fp32 [n, d] */
int echo_embedding(const int32_t* ids, int n, const float* table, int d, float* out, echo_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * S1-DAC decode helpers (time-major [B, T, C] activations).
 * ------------------------------------------------------------------------------------------ */
/* ConvNeXt front: causal depthwise conv k=7 + LayerNorm(eps) -> bf16 (autoencoder.py:362-364) */
int echo_dwconv_layernorm(const float* x, int batch, int t, int c, const float* w_dw /*[c,7]*/,
                          const float* b_dw, const float* ln_w, const float* ln_b, float eps,
                          void* y_bf16, echo_stream_t stream);
/* last stage: causal conv k=7, C -> 1, + tanh (autoencoder.py:994): x bf16 [B,T,C] (already Snake'd) */
int echo_conv_out_tanh(const void* x_bf16, int batch, int t, int c, const float* w /*[7][c]*/, float bias,
                       float* wav, echo_stream_t stream);
/* Streaming forms of the two causal k=7 kernels above (S1-DAC decode stream, batch 1): `left_ctx` (0..6) rows of the
 * previous call's input sit in memory immediately BEFORE x and are read where the one-shot kernels pad with zeros
 * (autoencoder.py:285-289); the first call of a stream passes an all-zero context = the reference's left padding. */
int echo_dwconv_layernorm_ctx(const float* x, int batch, int t, int c, int left_ctx, const float* w_dw,
                              const float* b_dw, const float* ln_w, const float* ln_b, float eps, void* y_bf16,
                              echo_stream_t stream);
int echo_conv_out_tanh_ctx(const void* x_bf16, int batch, int t, int c, int left_ctx, const float* w, float bias,
                           float* wav, echo_stream_t stream);
/* element-wise fp32 -> bf16 with optional Snake (autoencoder.py:96-102) over [rows, c] */
int echo_cast_snake(const float* x, int64_t rows, int c, const float* alpha, void* y_bf16, echo_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Conditioning encoders as ONE native call each (host code, csrc/encode_kv.cu): a TextEncoder / SpeakerEncoder stack
 * (EncoderTransformerBlock, model.py:311-339; SelfAttention :106-161) followed by the per-decoder-layer key/value
 * projections (JointAttention.get_kv_cache_text/speaker/latent, model.py:270-293).  Replaces
 * EchoDiT.get_kv_cache_text / _speaker / _latent (model.py:606-636) after the embedding / patch projection.
 * Incremental form (`past` > 0): the encoder is causal, so positions already encoded never change -- only the new
 * positions are run, against the cached per-layer encoder K/V, and their projections are appended (blockwise
 * generation re-encodes the whole prefix three times per block in the reference, inference_blockwise.py:72-74).
 * ------------------------------------------------------------------------------------------ */
typedef struct echo_enc_layer {
  const void* w_qkvg;       /* bf16 [4*d, d]  (wq | wk | wv | gate) */
  const void* w_o;          /* bf16 [d, d] */
  const void* w13;          /* bf16 [2*ffn, d], 64-row groups w1 | w3 */
  const void* w2;           /* bf16 [d, ffn] */
  const float *q_norm, *k_norm;        /* per-head RMSNorm weights [d] */
  const float *attn_norm, *mlp_norm;   /* RMSNorm [d] */
} echo_enc_layer;

typedef struct echo_encode_kv_desc {
  int32_t n_layers, d, heads, ffn;     /* encoder; head_dim is 128 */
  float eps;
  int32_t causal;                      /* 1: SpeakerEncoder / latent encoder, 0: TextEncoder */
  const echo_enc_layer* layers;        /* HOST array [n_layers] */
  const float* out_norm;               /* text_norm / speaker_norm / latent_norm [d] (model.py:608, 617, 628) */
  const float *rope_cos, *rope_sin;    /* fp32 [rope_len][64] */
  int32_t rope_len;
  int32_t rows, seq, past;             /* prompts, NEW positions per prompt, positions encoded by earlier calls */
  float* x;                            /* fp32 [rows*seq, d]: embedding / patch projection; clobbered (residual stream) */
  const uint8_t* mask;                 /* key-valid bytes [rows][mask_stride] or NULL (text padding mask, model.py:410-418) */
  int64_t mask_stride;
  const int32_t* extent;               /* per-row 1 + last valid key, or NULL */
  void* enc_cache;                     /* bf16 [n_layers][2][rows][enc_cap][d] encoder K / V, or NULL (then past must be 0) */
  int32_t enc_cap;
  int32_t dec_layers, dec_d, dec_heads;
  const void* const* w_kv;             /* HOST array [dec_layers]: bf16 [2*dec_d, d]  (wk_* | wv_*) */
  const float* const* k_norm;          /* HOST array [dec_layers]: fp32 [dec_d] */
  int32_t rope_heads, pos_stride;      /* latent prefix: dec_heads/2 heads rotated at position pos_stride * index (model.py:284-293); else 0, 1 */
  void* kv_out;                        /* bf16 [dec_layers][2][rows][out_cap][dec_d]; this call writes positions [past, past+seq) */
  int32_t out_cap, gemm_impl;
  void* gemm_workspace;
  int64_t gemm_workspace_bytes;
} echo_encode_kv_desc;
int64_t echo_encode_kv_workspace_bytes(const echo_encode_kv_desc* d);
int echo_encode_kv(const echo_encode_kv_desc* d, void* workspace, int64_t workspace_bytes, echo_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Whole S1-DAC decode (DAC.decode_zq, autoencoder.py:1128-1132) as ONE native call, and its incremental form.
 * Host code only (csrc/ae_decode.cu): a launch sequence over echo_gemm_bf16 / echo_attention / echo_rmsnorm_mod /
 * echo_dwconv_layernorm_ctx / echo_conv_out_tanh_ctx.  All pointers are device pointers to the caller's packed weights
 * (layouts: echo_tts_b200/autoencoder.py, the packing functions cite the reference modules); the descriptor arrays
 * themselves are HOST memory.  No allocation inside: the caller lends a workspace (echo_ae_decode_workspace_bytes).
 * ------------------------------------------------------------------------------------------ */
typedef struct echo_ae_tr_layer {      /* one TransformerBlock of post_module (autoencoder.py:682-724) */
  const void* wqkv;         /* bf16 [3*heads*head_dim, d] */
  const void* wo;           /* bf16 [d, heads*head_dim] */
  const void* w13;          /* bf16 [2*ffn, d], rows interleaved in groups of 64 (w1 | w3) */
  const void* w2;           /* bf16 [d, ffn] */
  const float *attn_norm, *ffn_norm;     /* RMSNorm weights [d] */
  const float *attn_scale, *ffn_scale;   /* LayerScale gamma [d] */
} echo_ae_tr_layer;

typedef struct echo_ae_upsample {      /* CausalTransConvNet (k == stride) + ConvNeXtBlock (autoencoder.py:427-435, 340-373) */
  int32_t factor, hidden;   /* upsampling factor; ConvNeXt hidden width */
  const void* wt;           /* bf16 [factor*d, d] (pixel-shuffle view of the transposed conv) */
  const float* bt;          /* [factor*d] */
  const float *dw, *dw_b;   /* depthwise k=7 [d, 7], [d] */
  const float *ln_w, *ln_b; /* LayerNorm [d] */
  const void* pw1;          /* bf16 [hidden, d] */
  const float* pb1;
  const void* pw2;          /* bf16 [d, hidden] */
  const float* pb2;
  const float* gamma;       /* [d] */
} echo_ae_upsample;

typedef struct echo_ae_unit {          /* ResidualUnit (autoencoder.py:879-900) */
  int32_t dilation, _pad;
  const float *alpha1, *alpha2;   /* Snake alphas [c] */
  const void* w7;           /* bf16 [c, 7*c] tap-major */
  const float* b7;
  const void* w1;           /* bf16 [c, c] */
  const float* b1;
} echo_ae_unit;

typedef struct echo_ae_block {         /* DecoderBlock (autoencoder.py:932-968) */
  int32_t rate, c_in, c_out, _pad;
  const float* alpha_in;    /* Snake before the transposed conv [c_in]: applied by the PRODUCER of this block's input */
  const float* alpha1_rep;  /* units[0].alpha1 repeated `rate` times [rate*c_out] */
  const void* wt;           /* bf16 [rate*c_out, 2*c_in] (taps 0, -1) */
  const float* bt;          /* [rate*c_out] */
  echo_ae_unit units[3];
} echo_ae_block;

typedef struct echo_ae_decoder_desc {
  int32_t d, tr_layers, tr_heads, tr_head_dim, tr_ffn, tr_window;
  float tr_eps, _pad0;
  const echo_ae_tr_layer* tr;             /* HOST array [tr_layers] */
  const float* tr_norm;                   /* final RMSNorm [d] */
  const float *rope_cos, *rope_sin;       /* fp32 [rope_len][tr_head_dim/2] (the module's own bf16-rounded table) */
  int32_t rope_len, n_up;
  const echo_ae_upsample* ups;            /* HOST array [n_up] */
  const void* conv0_w;                    /* bf16 [c0, 7*d] */
  const float* conv0_b;
  int32_t c0, n_blocks;
  const echo_ae_block* blocks;            /* HOST array [n_blocks] */
  const float* alpha_out;                 /* Snake before the last conv [c_last] */
  const float* w_last;                    /* fp32 [7][c_last] */
  float b_last;
  int32_t gemm_impl;
  void* gemm_workspace;                   /* optional stream-K scratch (echo_gemm_desc.workspace) */
  int64_t gemm_workspace_bytes;
} echo_ae_decoder_desc;

/* bytes of scratch echo_ae_decode (stream == 0) / echo_ae_decode_stream_step (stream != 0, batch 1) need for `frames` */
int64_t echo_ae_decode_workspace_bytes(const echo_ae_decoder_desc* d, int batch, int frames, int stream);
/* x: fp32 [batch*frames, d] = z_q time-major (after the PCA inverse, inference.py:234), CLOBBERED (it is the residual
 * stream of post_module); wav: fp32 [batch, frames*hop]. */
int echo_ae_decode(const echo_ae_decoder_desc* d, float* x, int batch, int frames, float* wav, void* workspace,
                   int64_t workspace_bytes, echo_stream_t stream);
/* Incremental decode of ONE utterance (no reference counterpart; README.md:124 leaves streaming out).  `state` carries,
 * per layer, what the next call needs: the last window-1 keys/values of every post_module layer and the last
 * (k-1)*dilation input rows of every causal conv (zeros = the reference's left padding, autoencoder.py:283-289).
 * `position` = frames decoded so far (RoPE position; caller-owned).  Results equal echo_ae_decode of the concatenation. */
int64_t echo_ae_decode_stream_state_bytes(const echo_ae_decoder_desc* d);
int echo_ae_decode_stream_init(const echo_ae_decoder_desc* d, void* state, int64_t state_bytes, echo_stream_t stream);
int echo_ae_decode_stream_step(const echo_ae_decoder_desc* d, void* state, float* x, int frames, int position,
                               float* wav, void* workspace, int64_t workspace_bytes, echo_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * S1-DAC encode helpers (DAC.encode_zq, autoencoder.py:1080-1126; the convolutions, transformers and ConvNeXt blocks
 * of the encode path run on echo_gemm_bf16 / echo_attention / echo_dwconv_layernorm like the decode path).
 * ------------------------------------------------------------------------------------------ */
/* Encoder's first layer (autoencoder.py:917) + the first ResidualUnit's Snake (:884): audio fp32 [B, T] ->
 * x fp32 [B, T, C] = causal conv k=7 (w [C][7]) + bias, and s bf16 [B, T, C] = snake(x, alpha) */
int echo_conv_in_snake(const float* audio, int batch, int t, int c, const float* w, const float* bias,
                       const float* alpha, float* x, void* s_bf16, echo_stream_t stream);
/* Snake1d on a bf16 operand (autoencoder.py:96-109): x bf16 [rows, c] -> y bf16 [rows, c] */
int echo_snake_bf16(const void* x_bf16, int64_t rows, int c, const float* alpha, void* y_bf16, echo_stream_t stream);
/* RMSNorm with an fp32 result (autoencoder.py:726-731), y = x * rsqrt(mean x^2 + eps) * weight: the quantizer input */
int echo_rmsnorm_f32(const float* x, int64_t ldx, int m, int d, float eps, const float* weight, float* y, int64_t ldy,
                     echo_stream_t stream);
/* VectorQuantize.decode_latents (autoencoder.py:145-157) in fp32: e [m, dim] (l2-normalised inside), codebook rows
 * given both l2-normalised (search) and raw (gather); index_out[m] = nearest code (lowest index on ties),
 * code_out (nullable) fp32 [m, dim] = the raw code vectors (embed_code :138-139).  dim must be 8. */
int echo_vq_nearest(const float* e, int m, int dim, const float* codebook_normalized, const float* codebook, int n,
                    int32_t* index_out, float* code_out, echo_stream_t stream);

/* fp32-class arithmetic of the encode path (csrc/precise.cu): the RVQ nearest-code search is an integer decision, so
 * every contraction in front of it (Encoder autoencoder.py:903-929, downsample :416-425, pre_module :744-802) runs on
 * echo_gemm_bf16 with both operands split in two bf16 terms and three products accumulated in fp32
 * (x w ~= hi w_hi + lo w_hi + hi w_lo): the activation operand is laid out [hi | lo | hi] along K by echo_split_bf16x3,
 * the weights are packed [w_hi | w_hi | w_lo] per tap at load.
 * echo_split_bf16x3: x fp32 [rows, c] (row stride ldx) -> out bf16 [rows, 3c] (row stride ldo).
 *   mode 0: y = x;  1: y = snake(x, alpha[c]) (Snake1d :96-109, precise sinf);
 *   mode 2: y = silu(x[:, :c]) * x[:, c:2c] (FeedForward :717; x has 2c columns). */
int echo_split_bf16x3(const float* x, int64_t ldx, int64_t rows, int c, int mode, const float* alpha, void* out_bf16,
                      int64_t ldo, echo_stream_t stream);
/* apply_rotary_emb (autoencoder.py:816-826) in place on fp32 rows [rows, ldx]: heads x head_dim columns, interleaved
 * pairs, position = pos_offset + row % rows_per_seq; cos / sin fp32 [rope_len, head_dim / 2]. */
int echo_rope_f32(float* x, int64_t ldx, int64_t rows, int heads, int head_dim, const float* cos_t, const float* sin_t,
                  int rope_len, int rows_per_seq, int pos_offset, echo_stream_t stream);
/* fp32 causal sliding-window attention (WindowLimitedTransformer mask, autoencoder.py:762-773: key j visible to query
 * i iff i - window < j <= i): q / k / v fp32 rows [batch * t, ld] with head h at columns h * 64 (head_dim must be 64),
 * out fp32 [batch * t, ldo]. */
int echo_attention_window_f32(const float* q, const float* k, const float* v, int64_t ld, int batch, int t, int heads,
                              int head_dim, int window, float softmax_scale, float* out, int64_t ldo, echo_stream_t stream);
/* echo_dwconv_layernorm with an fp32 result (ConvNeXtBlock front, autoencoder.py:362-364). */
int echo_dwconv_layernorm_f32(const float* x, int batch, int t, int c, const float* w_dw, const float* b_dw,
                              const float* ln_w, const float* ln_b, float eps, float* y, echo_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Audio front-end (load_audio, inference.py:141-149), after the file has been decoded to PCM on the host.
 * ------------------------------------------------------------------------------------------ */
/* Channel mean (inference.py:145) + polyphase windowed-sinc resampling = torchaudio.functional.resample (inference.py:146;
 * third party: torchaudio 2.x `_apply_sinc_resample_kernel`, restated in oracle/echo_oracle.py::resample_sinc):
 *   y[f*up + p] = sum_k table_t[k*up + p] * mono[f*down + k - width],   k in [0, 2*width + down), zeros outside [0, n)
 * x fp32 [channels][ldx], (down, up) = (orig, new) / gcd, table_t fp32 [2*width + down][up] (host-built filter bank,
 * transposed), y fp32 [n_out] with n_out = ceil(up * n / down).  absmax_bits (nullable, zeroed by the caller) receives
 * the bit pattern of max |y| (inference.py:147's peak search fused into the same pass).
 * tap_lo / tap_hi (nullable, int32 [up]): per-phase range [lo, hi) of taps whose weight is not negligible -- the Hann
 * window is zero beyond +-lowpass_filter_width zero crossings, so ~2*width of the 2*width + down taps matter; the caller
 * derives the ranges from the table (torchaudio multiplies through all of them). */
int echo_resample_mono(const float* x, int channels, int64_t n, int64_t ldx, int down, int up, int width,
                       const float* table_t, const int32_t* tap_lo, const int32_t* tap_hi, float* y, int64_t n_out,
                       uint32_t* absmax_bits, echo_stream_t stream);
/* inference.py:147: y /= max(peak, floor) with peak read from absmax_bits (floor = 1: never amplify) */
int echo_peak_normalize(float* y, int64_t n, const uint32_t* absmax_bits, float floor_, echo_stream_t stream);

/* find_flattening_point (inference.py:315-330): latent fp32 [len, width] -> index (device int32) */
int echo_flattening_point(const float* latent, int len, int width, int window, float std_threshold,
                          float target, int32_t* index_out, echo_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* ECHO_B200_H */
