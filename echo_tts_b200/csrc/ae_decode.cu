// Whole S1-DAC decode as one native call, one-shot and incremental (host code only): a launch sequence over the kernels
// of this library.  Reference: DAC.decode_zq autoencoder.py:1128-1132 = post_module (WindowLimitedTransformer,
// :744-802) -> quantizer.upsample (:427-435, ConvNeXtBlock :340-373) -> Decoder (:971-998, DecoderBlock :932-968,
// ResidualUnit :879-900).  Activations are time-major [B, T, C]; every conv is an implicit GEMM over shifted row
// windows.  In stream mode (batch 1) each conv input is laid out [context rows | new rows] and the GEMM is pointed at
// the new rows with a_ctx context rows in front; the context rows are carried in `state` between calls.
#include <string.h>
#include "common.cuh"
#include "../../include/echo_b200.h"

namespace {

constexpr int LEAD = 54;    // rows in front of the shared Snake buffer: the widest context, (7 - 1) * dilation 9

inline int64_t align_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

// Two arenas used alternately by consecutive stages: a stage reads only what the previous one produced.
struct Arena {
  uint8_t* base = nullptr;
  int64_t half = 0;          // bytes per arena (0 in a dry run)
  int64_t off[2] = {0, 0}, peak[2] = {0, 0};
  int cur = 1;
  bool dry = true;
  void next_stage() { cur ^= 1; off[cur] = 0; }
  template <typename T>
  T* take(int64_t elems) {
    const int64_t bytes = align_up(elems * (int64_t)sizeof(T), 256);
    const int64_t o = off[cur];
    off[cur] += bytes;
    if (off[cur] > peak[cur]) peak[cur] = off[cur];
    return dry ? nullptr : reinterpret_cast<T*>(base + (int64_t)cur * half + o);
  }
};

// Offsets (bytes) of the carried rows inside the stream state, in the order the decode walks the layers.
struct StateCursor {
  uint8_t* base;
  int64_t off = 0;
  template <typename T>
  T* take(int64_t elems) {
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += align_up(elems * (int64_t)sizeof(T), 256);
    return p;
  }
};

struct Ctx {
  const echo_ae_decoder_desc* D;
  cudaStream_t st;
  Arena ws;
  StateCursor state;
  bool stream;               // incremental mode (batch 1, carried context)
  int pos;                   // frames decoded before this call
  int rc = ECHO_OK;
  bool live() const { return !ws.dry && rc == ECHO_OK; }
  void run(int r) { if (rc == ECHO_OK && r != ECHO_OK) rc = r; }
  void copy(void* dst, const void* src, int64_t bytes) {
    if (live() && bytes > 0) {
      if (cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToDevice, st) != cudaSuccess) {
        echo_set_error("ae_decode: state copy failed: %s", cudaGetErrorString(cudaGetLastError()));
        rc = ECHO_ERR_CUDA;
      }
    }
  }
  // carried rows -> buf[lead - n, lead); afterwards keep the last n rows of buf[0, lead + t)
  template <typename T>
  void load(T* carried, T* buf, int lead, int n, int c) {
    if (live()) copy(buf + (int64_t)(lead - n) * c, carried, (int64_t)n * c * sizeof(T));
  }
  template <typename T>
  void save(T* carried, const T* buf, int lead, int t, int n, int c) {
    if (live()) copy(carried, buf + (int64_t)(lead + t - n) * c, (int64_t)n * c * sizeof(T));
  }
};

void gemm_init(const Ctx& c, echo_gemm_desc* g, const void* a, int64_t lda, int rows, int batch, int k_tap, const void* w,
               int n, int ntaps, const int* taps, int a_ctx, int epilogue) {
  memset(g, 0, sizeof(*g));
  g->a = a; g->b = w;
  g->lda = lda; g->ldb = (int64_t)ntaps * k_tap;
  g->rows = rows; g->batch = batch; g->a_batch_stride = (int64_t)rows * lda;
  g->n = n; g->k_tap = k_tap; g->taps = ntaps;
  for (int i = 0; i < ntaps; ++i) g->tap_shift[i] = taps ? taps[i] : 0;
  g->a_ctx = a_ctx;
  g->epilogue = epilogue;
  g->impl = c.D->gemm_impl;
  g->rows_per_gate = 1;
  g->b_static = 1;             // w is a load-time tensor of the descriptor (weights)
  if (c.D->gemm_impl >= 2) { g->workspace = c.D->gemm_workspace; g->workspace_bytes = c.D->gemm_workspace_bytes; }
}

// x += gamma * (a w^T)   (LayerScale residual, autoencoder.py:718-723)
void scaled_residual(Ctx& c, const void* a, int k, const void* w, const float* gamma, float* x, int m, int d) {
  if (!c.live()) return;
  echo_gemm_desc g;
  gemm_init(c, &g, a, k, m, 1, k, w, d, 1, nullptr, 0, ECHO_EPI_GENERIC);
  g.gate = gamma; g.gate_ld = 0; g.rows_per_gate = 1 << 30;
  g.resid = x; g.ld_resid = d; g.out_f32 = x; g.ld_out_f32 = d;
  c.run(echo_gemm_bf16(&g, c.st));
}

void rmsnorm(Ctx& c, const float* x, int m, int d, const float* w, void* y) {
  if (!c.live()) return;
  c.run(echo_rmsnorm_mod(x, d, m, d, c.D->tr_eps, w, nullptr, 0, 0, nullptr, 0, y, d, c.st));
}

// post_module: x fp32 [B*T, d] residual stream (in place) -> final RMSNorm as bf16 [B*T, d]
bf16* transformer(Ctx& c, float* x, int B, int T) {
  const echo_ae_decoder_desc* D = c.D;
  const int d = D->d, H = D->tr_heads, Dh = D->tr_head_dim, HD = H * Dh, W = D->tr_window;
  const int M = B * T;
  const int C1 = c.stream ? W - 1 : 0;                 // cached keys a new query can still see
  const int nv = c.stream ? (c.pos < C1 ? c.pos : C1) : 0;
  c.ws.next_stage();
  bf16* xn = c.ws.take<bf16>((int64_t)M * d);
  bf16* q = c.ws.take<bf16>((int64_t)M * HD);
  bf16* kbuf = c.ws.take<bf16>((int64_t)(C1 + T) * B * HD);
  bf16* vbuf = c.ws.take<bf16>((int64_t)(C1 + T) * B * HD);
  bf16* ao = c.ws.take<bf16>((int64_t)M * HD);
  bf16* hmid = c.ws.take<bf16>((int64_t)M * D->tr_ffn);
  for (int l = 0; l < D->tr_layers; ++l) {
    const echo_ae_tr_layer& lw = D->tr[l];
    bf16* kc = c.stream ? c.state.take<bf16>((int64_t)C1 * HD) : nullptr;
    bf16* vc = c.stream ? c.state.take<bf16>((int64_t)C1 * HD) : nullptr;
    rmsnorm(c, x, M, d, lw.attn_norm, xn);
    if (c.stream) { c.load(kc, kbuf, C1, C1, HD); c.load(vc, vbuf, C1, C1, HD); }
    if (c.live()) {
      echo_gemm_desc g;
      gemm_init(c, &g, xn, d, M, 1, d, lw.wqkv, 3 * HD, 1, nullptr, 0, ECHO_EPI_SECTIONS);
      g.n_sections = 3; g.section_width = HD; g.head_dim = Dh; g.rope_heads = H;
      g.pos_offset = c.pos; g.pos_stride = 1; g.rows_per_seq = T; g.eps = D->tr_eps;
      g.rope_cos = D->rope_cos; g.rope_sin = D->rope_sin; g.ld_sec_out = HD;
      g.sections[0].type = ECHO_SEC_ROPE; g.sections[0].out = q;
      g.sections[1].type = ECHO_SEC_ROPE; g.sections[1].out = kbuf + (int64_t)C1 * HD;
      g.sections[2].type = ECHO_SEC_PLAIN; g.sections[2].out = vbuf + (int64_t)C1 * HD;
      c.run(echo_gemm_bf16(&g, c.st));
    }
    if (c.live()) {
      echo_attn_desc a;
      memset(&a, 0, sizeof(a));
      a.q = q; a.q_row_stride = HD; a.q_batch_stride = (int64_t)T * HD;
      a.out = ao;
      a.rows = B; a.seq = T; a.heads = H; a.head_dim = Dh;
      a.n_seg = 1; a.impl = 0;
      a.softmax_scale = 1.0f / sqrtf((float)Dh);
      echo_attn_segment& s = a.seg[0];
      s.k = kbuf + (int64_t)(C1 - nv) * HD; s.v = vbuf + (int64_t)(C1 - nv) * HD;
      s.row_stride = HD; s.batch_stride = (int64_t)(C1 + T) * HD;
      s.len = nv + T; s.causal = 1; s.window = W; s.q_offset = nv;
      c.run(echo_attention(&a, c.st));
    }
    if (c.stream) { c.save(kc, kbuf, C1, T, C1, HD); c.save(vc, vbuf, C1, T, C1, HD); }
    scaled_residual(c, ao, HD, lw.wo, lw.attn_scale, x, M, d);
    rmsnorm(c, x, M, d, lw.ffn_norm, xn);
    if (c.live()) {
      echo_gemm_desc g;
      gemm_init(c, &g, xn, d, M, 1, d, lw.w13, 2 * D->tr_ffn, 1, nullptr, 0, ECHO_EPI_SWIGLU);
      g.out_bf16 = hmid; g.ld_out_bf16 = D->tr_ffn;
      c.run(echo_gemm_bf16(&g, c.st));
    }
    scaled_residual(c, hmid, D->tr_ffn, lw.w2, lw.ffn_scale, x, M, d);
  }
  rmsnorm(c, x, M, d, D->tr_norm, xn);
  return xn;
}

// quantizer.upsample: a bf16 [B*T, d] -> bf16 buffer whose rows lead6.. are the [B*T', d] result (lead6 = 6 in stream
// mode: conv0's context lives in front of it)
bf16* upsample(Ctx& c, const bf16* a, int B, int& T) {
  const echo_ae_decoder_desc* D = c.D;
  const int d = D->d, lead = c.stream ? 6 : 0;
  bf16* out = nullptr;
  for (int u = 0; u < D->n_up; ++u) {
    const echo_ae_upsample& up = D->ups[u];
    const int fac = up.factor, M = B * T;
    c.ws.next_stage();
    float* xb = c.ws.take<float>((int64_t)(lead + (int64_t)M * fac) * d);
    float* x = xb ? xb + (int64_t)lead * d : nullptr;
    if (c.live()) {   // transposed conv with k == stride: one tap, N = fac*d (pixel-shuffle view), no trim
      echo_gemm_desc g;
      gemm_init(c, &g, a, d, M, 1, d, up.wt, fac * d, 1, nullptr, 0, ECHO_EPI_GENERIC);
      g.bias = up.bt; g.out_f32 = x; g.ld_out_f32 = (int64_t)fac * d;
      c.run(echo_gemm_bf16(&g, c.st));
    }
    T *= fac;
    const int M2 = B * T;
    float* dwc = c.stream ? c.state.take<float>(6 * (int64_t)d) : nullptr;
    if (c.stream) c.load(dwc, xb, 6, 6, d);
    bf16* y = c.ws.take<bf16>((int64_t)M2 * d);
    if (c.live())
      c.run(echo_dwconv_layernorm_ctx(x, B, T, d, lead, up.dw, up.dw_b, up.ln_w, up.ln_b, 1e-6f, y, c.st));
    if (c.stream) c.save(dwc, xb, 6, T, 6, d);
    bf16* hmid = c.ws.take<bf16>((int64_t)M2 * up.hidden);
    if (c.live()) {
      echo_gemm_desc g;
      gemm_init(c, &g, y, d, M2, 1, d, up.pw1, up.hidden, 1, nullptr, 0, ECHO_EPI_GENERIC);
      g.bias = up.pb1; g.act = ECHO_ACT_GELU; g.out_bf16 = hmid; g.ld_out_bf16 = up.hidden;
      c.run(echo_gemm_bf16(&g, c.st));
    }
    out = c.ws.take<bf16>((int64_t)(lead + M2) * d);
    bf16* an = out ? out + (int64_t)lead * d : nullptr;
    if (c.live()) {
      echo_gemm_desc g;
      gemm_init(c, &g, hmid, up.hidden, M2, 1, up.hidden, up.pw2, d, 1, nullptr, 0, ECHO_EPI_GENERIC);
      g.bias = up.pb2; g.gate = up.gamma; g.gate_ld = 0; g.rows_per_gate = 1 << 30;
      g.resid = x; g.ld_resid = d; g.out_bf16 = an; g.ld_out_bf16 = d;
      c.run(echo_gemm_bf16(&g, c.st));
    }
    a = an;
  }
  return out;
}

// Decoder: ab = bf16 buffer whose rows lead6.. are [B*T, d]  ->  wav fp32 [B, T * prod(rates)]
void wave_decoder(Ctx& c, bf16* ab, int B, int T, float* wav) {
  const echo_ae_decoder_desc* D = c.D;
  const int d = D->d;
  const int lead6 = c.stream ? 6 : 0, lead1 = c.stream ? 1 : 0, leadU = c.stream ? LEAD : 0;
  int taps7[7];
  for (int k = 0; k < 7; ++k) taps7[k] = -(6 - k);
  c.ws.next_stage();
  bf16* c0c = c.stream ? c.state.take<bf16>(6 * (int64_t)d) : nullptr;
  if (c.stream) c.load(c0c, ab, 6, 6, d);
  bf16* sb = c.ws.take<bf16>((int64_t)(lead1 + (int64_t)B * T) * D->c0);
  if (c.live()) {
    echo_gemm_desc g;
    gemm_init(c, &g, ab + (int64_t)lead6 * d, d, T, B, d, D->conv0_w, D->c0, 7, taps7, lead6, ECHO_EPI_GENERIC);
    g.bias = D->conv0_b; g.out2_bf16 = sb + (int64_t)lead1 * D->c0; g.ld_out2_bf16 = D->c0;
    g.snake_alpha = D->blocks[0].alpha_in;
    c.run(echo_gemm_bf16(&g, c.st));
  }
  if (c.stream) c.save(c0c, ab, 6, T, 6, d);
  int lead = lead1;
  int c_last = D->c0;
  for (int bi = 0; bi < D->n_blocks; ++bi) {
    const echo_ae_block& blk = D->blocks[bi];
    const int r = blk.rate, cin = blk.c_in, co = blk.c_out;
    c.ws.next_stage();
    float* x = c.ws.take<float>((int64_t)B * T * r * co);
    bf16* s1b = c.ws.take<bf16>((int64_t)(leadU + (int64_t)B * T * r) * co);
    bf16* s1 = s1b ? s1b + (int64_t)leadU * co : nullptr;
    bf16* upc = c.stream ? c.state.take<bf16>(cin) : nullptr;
    if (c.stream) c.load(upc, sb, lead, 1, cin);
    if (c.live()) {   // transposed conv k = 2*stride: taps (0, -1), N = r*c_out, + the first unit's Snake
      const int tp[2] = {0, -1};
      echo_gemm_desc g;
      gemm_init(c, &g, sb + (int64_t)lead * cin, cin, T, B, cin, blk.wt, r * co, 2, tp, c.stream ? 1 : 0, ECHO_EPI_GENERIC);
      g.bias = blk.bt; g.out_f32 = x; g.ld_out_f32 = (int64_t)r * co;
      g.out2_bf16 = s1; g.ld_out2_bf16 = (int64_t)r * co; g.snake_alpha = blk.alpha1_rep;
      c.run(echo_gemm_bf16(&g, c.st));
    }
    if (c.stream) c.save(upc, sb, lead, T, 1, cin);
    T *= r;
    bf16* s2 = c.ws.take<bf16>((int64_t)B * T * co);
    for (int j = 0; j < 3; ++j) {
      const echo_ae_unit& u = blk.units[j];
      const int dil = u.dilation, n = 6 * dil;
      int tp[7];
      for (int k = 0; k < 7; ++k) tp[k] = -(6 - k) * dil;
      bf16* uc = c.stream ? c.state.take<bf16>((int64_t)n * co) : nullptr;
      if (c.stream) c.load(uc, s1b, LEAD, n, co);
      if (c.live()) {
        echo_gemm_desc g;
        gemm_init(c, &g, s1, co, T, B, co, u.w7, co, 7, tp, c.stream ? n : 0, ECHO_EPI_GENERIC);
        g.bias = u.b7; g.out2_bf16 = s2; g.ld_out2_bf16 = co; g.snake_alpha = u.alpha2;
        c.run(echo_gemm_bf16(&g, c.st));
      }
      if (c.stream) c.save(uc, s1b, LEAD, T, n, co);
      const float* nxt = j + 1 < 3 ? blk.units[j + 1].alpha1 : (bi + 1 < D->n_blocks ? D->blocks[bi + 1].alpha_in : D->alpha_out);
      if (c.live()) {
        echo_gemm_desc g;
        gemm_init(c, &g, s2, co, B * T, 1, co, u.w1, co, 1, nullptr, 0, ECHO_EPI_GENERIC);
        g.bias = u.b1; g.resid = x; g.ld_resid = co;
        if (j + 1 < 3) { g.out_f32 = x; g.ld_out_f32 = co; }     // after the last unit only snake(x) is read again
        g.out2_bf16 = s1; g.ld_out2_bf16 = co; g.snake_alpha = nxt;
        c.run(echo_gemm_bf16(&g, c.st));
      }
    }
    sb = s1b; lead = leadU; c_last = co;
  }
  bf16* oc = c.stream ? c.state.take<bf16>(6 * (int64_t)c_last) : nullptr;
  if (c.stream) c.load(oc, sb, lead, 6, c_last);
  if (c.live())
    c.run(echo_conv_out_tanh_ctx(sb + (int64_t)lead * c_last, B, T, c_last, c.stream ? 6 : 0, D->w_last, D->b_last, wav, c.st));
  if (c.stream) c.save(oc, sb, lead, T, 6, c_last);
}

int check_desc(const echo_ae_decoder_desc* D) {
  ECHO_CHECK_ARG(D != nullptr, "ae_decode: null descriptor");
  ECHO_CHECK_ARG(D->d > 0 && D->tr_layers >= 1 && D->n_up >= 1 && D->tr_heads > 0 && (D->tr_head_dim == 64 || D->tr_head_dim == 128),
                 "ae_decode: bad transformer shape");
  ECHO_CHECK_ARG(D->tr && D->tr_norm && D->rope_cos && D->rope_sin, "ae_decode: null transformer weights");
  ECHO_CHECK_ARG(D->tr_window > 1 && D->tr_ffn % 64 == 0, "ae_decode: bad window / ffn");
  ECHO_CHECK_ARG(D->ups != nullptr, "ae_decode: null upsample array");
  ECHO_CHECK_ARG(D->n_blocks > 0 && D->blocks && D->conv0_w && D->conv0_b && D->alpha_out && D->w_last,
                 "ae_decode: null decoder weights");
  for (int i = 0; i < D->n_blocks; ++i)
    for (int j = 0; j < 3; ++j)
      ECHO_CHECK_ARG(D->blocks[i].units[j].dilation >= 1 && 6 * D->blocks[i].units[j].dilation <= LEAD,
                     "ae_decode: dilation %d not supported (<= 9)", D->blocks[i].units[j].dilation);
  return ECHO_OK;
}

// One walk of the decode: dry (sizes only) when workspace == nullptr.
int walk(const echo_ae_decoder_desc* D, float* x, int B, int T, int pos, bool stream, void* state, float* wav,
         void* workspace, int64_t workspace_bytes, cudaStream_t st, int64_t* need_ws, int64_t* need_state) {
  Ctx c;
  c.D = D; c.st = st; c.stream = stream; c.pos = pos;
  c.ws.dry = workspace == nullptr;
  c.state.base = static_cast<uint8_t*>(state);
  if (!c.ws.dry) {
    c.ws.base = static_cast<uint8_t*>(workspace);
    c.ws.half = workspace_bytes / 2 / 256 * 256;
  }
  bf16* a = transformer(c, x, B, T);
  int T2 = T;
  bf16* ab = upsample(c, a, B, T2);
  wave_decoder(c, ab, B, T2, wav);
  const int64_t peak = c.ws.peak[0] > c.ws.peak[1] ? c.ws.peak[0] : c.ws.peak[1];
  if (need_ws) *need_ws = 2 * peak + 512;
  if (need_state) *need_state = c.state.off;
  if (!c.ws.dry) ECHO_CHECK_ARG(peak <= c.ws.half, "ae_decode: workspace too small (%lld bytes, need %lld)",
                                (long long)workspace_bytes, (long long)(2 * peak + 512));
  return c.rc;
}

}  // namespace

extern "C" int64_t echo_ae_decode_workspace_bytes(const echo_ae_decoder_desc* D, int batch, int frames, int stream) {
  if (check_desc(D) != ECHO_OK || batch <= 0 || frames <= 0 || (stream && batch != 1)) return -1;
  int64_t need = 0;
  walk(D, nullptr, batch, frames, 0, stream != 0, nullptr, nullptr, nullptr, 0, nullptr, &need, nullptr);
  return need;
}

extern "C" int echo_ae_decode(const echo_ae_decoder_desc* D, float* x, int batch, int frames, float* wav, void* workspace,
                              int64_t workspace_bytes, echo_stream_t stream) {
  if (int rc = check_desc(D)) return rc;
  ECHO_CHECK_ARG(x && wav && workspace && batch > 0 && frames > 0, "ae_decode: bad argument");
  ECHO_CHECK_ARG(D->tr_layers == 0 || frames <= D->rope_len, "ae_decode: %d frames exceed the RoPE table (%d)", frames, D->rope_len);
  ECHO_CHECK_ARG(((uintptr_t)workspace % 256) == 0, "ae_decode: workspace must be 256-byte aligned");
  int64_t need = 0;
  walk(D, nullptr, batch, frames, 0, false, nullptr, nullptr, nullptr, 0, nullptr, &need, nullptr);
  ECHO_CHECK_ARG(workspace_bytes >= need, "ae_decode: workspace too small (%lld bytes, need %lld)", (long long)workspace_bytes, (long long)need);
  return walk(D, x, batch, frames, 0, false, nullptr, wav, workspace, workspace_bytes, (cudaStream_t)stream, nullptr, nullptr);
}

extern "C" int64_t echo_ae_decode_stream_state_bytes(const echo_ae_decoder_desc* D) {
  if (check_desc(D) != ECHO_OK) return -1;
  int64_t need = 0;
  walk(D, nullptr, 1, 1, 0, true, nullptr, nullptr, nullptr, 0, nullptr, nullptr, &need);
  return need;
}

extern "C" int echo_ae_decode_stream_init(const echo_ae_decoder_desc* D, void* state, int64_t state_bytes,
                                          echo_stream_t stream) {
  if (int rc = check_desc(D)) return rc;
  const int64_t need = echo_ae_decode_stream_state_bytes(D);
  ECHO_CHECK_ARG(state && state_bytes >= need, "ae_decode_stream_init: state too small (%lld bytes, need %lld)", (long long)state_bytes, (long long)need);
  ECHO_CHECK_ARG(((uintptr_t)state % 256) == 0, "ae_decode_stream_init: state must be 256-byte aligned");
  ECHO_CUDA(cudaMemsetAsync(state, 0, (size_t)need, (cudaStream_t)stream));   // zero context = the reference's zero padding
  return ECHO_OK;
}

extern "C" int echo_ae_decode_stream_step(const echo_ae_decoder_desc* D, void* state, float* x, int frames, int position,
                                          float* wav, void* workspace, int64_t workspace_bytes, echo_stream_t stream) {
  if (int rc = check_desc(D)) return rc;
  ECHO_CHECK_ARG(state && x && wav && workspace && frames > 0 && position >= 0, "ae_decode_stream_step: bad argument");
  ECHO_CHECK_ARG(D->tr_layers == 0 || position + frames <= D->rope_len,
                 "ae_decode_stream_step: stream longer than the post_module RoPE table (%d)", D->rope_len);
  ECHO_CHECK_ARG(((uintptr_t)workspace % 256) == 0 && ((uintptr_t)state % 256) == 0, "ae_decode_stream_step: 256-byte alignment");
  int64_t need = 0;
  walk(D, nullptr, 1, frames, position, true, nullptr, nullptr, nullptr, 0, nullptr, &need, nullptr);
  ECHO_CHECK_ARG(workspace_bytes >= need, "ae_decode_stream_step: workspace too small (%lld bytes, need %lld)", (long long)workspace_bytes, (long long)need);
  return walk(D, x, 1, frames, position, true, state, wav, workspace, workspace_bytes, (cudaStream_t)stream, nullptr, nullptr);
}
