// Whole-step entry points (host code only): one DiT velocity evaluation and the Euler loop as launch sequences over the
// kernels of this library.  Reference: EchoDiT.forward model.py:563-604, TransformerBlock.forward model.py:371-390,
// JointAttention.forward model.py:204-268, MLP.forward model.py:296-308, sampling loop inference.py:508-558.
#include <math.h>
#include "common.cuh"
#include "../../include/echo_b200.h"

namespace {

struct StepShape {
  int M, d, S, R;
  int rows_per_mod;
  int64_t mod_ld;
};

// address of modulation vector `which` ({0 shift, 1 1+scale, 2 gate}) of norm `j` of layer `l`, table row 0
inline const float* mod_vec(const echo_dit_step_desc* D, int l, int j, int which) {
  return D->mod + ((int64_t)(l * 2 + j) * 3 + which) * D->d;
}

void gemm_base(const echo_dit_step_desc* D, echo_gemm_desc* g, const void* a, int64_t lda, int rows, int k,
               const void* w, int n, int epilogue) {
  memset(g, 0, sizeof(*g));
  g->a = a; g->b = w;
  g->lda = lda; g->ldb = k;
  g->rows = rows; g->batch = 1; g->a_batch_stride = (int64_t)rows * lda;
  g->n = n; g->k_tap = k; g->taps = 1;
  g->epilogue = epilogue;
  g->impl = D->gemm_impl;
  g->rows_per_gate = 1;
  g->b_static = 1;             // w is a load-time tensor of the descriptor (weights)
  if (D->gemm_impl >= 2) { g->workspace = D->gemm_workspace; g->workspace_bytes = D->gemm_workspace_bytes; }
}

// gated residual projection (model.py:385, 388): x += gate * (a w^T), or its split-K form into `delta`
int residual_projection(const echo_dit_step_desc* D, const StepShape& sh, const void* a, int k, const void* w, int l,
                        int j, cudaStream_t st) {
  echo_gemm_desc g;
  gemm_base(D, &g, a, k, sh.M, k, w, sh.d, ECHO_EPI_GENERIC);
  g.gate = mod_vec(D, l, j, 2);
  g.gate_ld = sh.mod_ld;
  g.rows_per_gate = sh.rows_per_mod > 0 ? sh.rows_per_mod : (1 << 30);
  if (D->k_splits > 1) {
    g.k_splits = D->k_splits;
    g.split_stride = (int64_t)sh.M * sh.d;
    g.out_f32 = D->delta; g.ld_out_f32 = sh.d;
  } else {
    g.resid = D->x; g.ld_resid = sh.d;
    g.out_f32 = D->x; g.ld_out_f32 = sh.d;
  }
  return echo_gemm_bf16(&g, st);
}

// AdaLN normalise + modulate (model.py:76-79) -> bf16 GEMM operand; folds pending split-K partials into x first
int adaln_norm(const echo_dit_step_desc* D, const StepShape& sh, int l, int j, bool pending, cudaStream_t st) {
  const float* scale1 = mod_vec(D, l, j, 1);
  const float* shift = mod_vec(D, l, j, 0);
  if (pending)
    return echo_add_rmsnorm_mod(D->x, sh.d, sh.M, sh.d, D->eps, D->delta, D->k_splits, (int64_t)sh.M * sh.d, scale1,
                                shift, sh.mod_ld, sh.rows_per_mod, D->xn, sh.d, st);
  return echo_rmsnorm_mod(D->x, sh.d, sh.M, sh.d, D->eps, scale1, shift, sh.mod_ld, sh.rows_per_mod, nullptr, 0, D->xn,
                          sh.d, st);
}

void kv_segment(echo_attn_segment* s, const echo_dit_kv& c, int l) {
  memset(s, 0, sizeof(*s));
  s->len = c.len;
  if (c.len > 0) {
    s->k = c.k[l]; s->v = c.v[l];
    s->row_stride = c.row_stride; s->batch_stride = c.batch_stride;
  }
  s->kv_batch_mod = c.kv_batch_mod;
  s->mask = c.mask; s->mask_stride = c.mask_stride;
  s->extent = c.extent;
  s->kv_scale = c.kv_scale ? c.kv_scale + l : nullptr;
}

int check_kv(const echo_dit_kv& c, const char* name, int rows) {
  ECHO_CHECK_ARG(c.len >= 0, "echo_dit_step: %s.len < 0", name);
  if (c.len > 0) {
    ECHO_CHECK_ARG(c.k && c.v, "echo_dit_step: %s: null k / v pointer arrays", name);
    ECHO_CHECK_ARG(c.kv_batch_mod >= 0 && c.kv_batch_mod <= rows, "echo_dit_step: %s.kv_batch_mod out of range", name);
  }
  return ECHO_OK;
}

}  // namespace

extern "C" int echo_dit_step(const echo_dit_step_desc* D, echo_stream_t stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ECHO_CHECK_ARG(D != nullptr, "echo_dit_step: null descriptor");
  ECHO_CHECK_ARG(D->n_layers > 0 && D->layers, "echo_dit_step: no layers");
  ECHO_CHECK_ARG(D->d > 0 && D->heads > 0 && D->d == D->heads * D->head_dim, "echo_dit_step: d != heads * head_dim");
  ECHO_CHECK_ARG(D->ffn > 0 && D->ffn % 64 == 0 && D->latent > 0, "echo_dit_step: bad ffn / latent size");
  ECHO_CHECK_ARG(D->rows > 0 && D->seq > 0 && D->x_rows > 0 && D->rows % D->x_rows == 0,
                 "echo_dit_step: rows (%d) must be a multiple of x_rows (%d)", D->rows, D->x_rows);
  ECHO_CHECK_ARG((int64_t)D->rows * D->seq < (1LL << 31), "echo_dit_step: too many tokens");
  ECHO_CHECK_ARG(D->mod && (D->mod_rows == 1 || D->mod_rows == D->rows), "echo_dit_step: mod_rows must be 1 or rows");
  ECHO_CHECK_ARG(D->x_in && D->in_w && D->in_b && D->x && D->xn && D->qkvg && D->ao && D->hmid,
                 "echo_dit_step: null input / weight / workspace pointer");
  ECHO_CHECK_ARG(D->rope_cos && D->rope_sin && D->pos_offset >= 0 && D->pos_offset + D->seq <= D->rope_len,
                 "echo_dit_step: rope table too short (%d positions, need %d)", D->rope_len, D->pos_offset + D->seq);
  ECHO_CHECK_ARG(D->k_splits <= 1 || D->delta, "echo_dit_step: k_splits > 1 needs delta");
  if (int rc = check_kv(D->prefix, "prefix", D->rows)) return rc;
  if (int rc = check_kv(D->text, "text", D->rows)) return rc;
  if (int rc = check_kv(D->speaker, "speaker", D->rows)) return rc;

  StepShape sh;
  sh.R = D->rows; sh.S = D->seq; sh.d = D->d; sh.M = D->rows * D->seq;
  sh.rows_per_mod = D->mod_rows > 1 ? D->seq : 0;
  sh.mod_ld = D->mod_row_stride;
  const int M = sh.M, d = sh.d;
  const int64_t Md = (int64_t)M * d;
  bf16* qkvg = static_cast<bf16*>(D->qkvg);

  {  // in_proj in fp32 (model.py:586); CFG rows: evaluated from the shared x_t, written to every branch
    const int reps = D->rows / D->x_rows, m = D->x_rows * D->seq;
    echo_sgemm_desc g;
    memset(&g, 0, sizeof(g));
    g.a = D->x_in; g.lda = D->latent;
    g.w = D->in_w; g.ldw = D->latent;
    g.bias = D->in_b;
    g.out = D->x; g.ldo = d; g.out_zstride = reps > 1 ? (int64_t)m * d : 0;
    g.m = m; g.n = d; g.k = D->latent; g.nz = reps;
    g.post = ECHO_POST_NONE; g.out_scale = 1.f;
    if (int rc = echo_sgemm_f32(&g, st)) return rc;
  }

  bool pending = false;   // split-K partial products of the last w2 not yet added to x
  for (int l = 0; l < D->n_layers; ++l) {
    const echo_dit_layer& W = D->layers[l];
    ECHO_CHECK_ARG(W.w_qkvg && W.w_o && W.w13 && W.w2 && W.q_norm && W.k_norm, "echo_dit_step: layer %d: null weight", l);
    if (int rc = adaln_norm(D, sh, l, 0, pending, st)) return rc;
    {  // q | k | v | gate projection, per-head RMSNorm + half-head RoPE on q, k, sigmoid on the gate (model.py:215-232)
      echo_gemm_desc g;
      gemm_base(D, &g, D->xn, d, M, d, W.w_qkvg, 4 * d, ECHO_EPI_SECTIONS);
      g.n_sections = 4; g.section_width = d; g.head_dim = D->head_dim; g.rope_heads = D->heads / 2;
      g.pos_offset = D->pos_offset; g.pos_stride = 1; g.rows_per_seq = D->seq; g.eps = D->eps;
      g.rope_cos = D->rope_cos; g.rope_sin = D->rope_sin;
      g.ld_sec_out = d;
      const int types[4] = {ECHO_SEC_NORM_ROPE, ECHO_SEC_NORM_ROPE, ECHO_SEC_PLAIN, ECHO_SEC_SIGMOID};
      const float* nw[4] = {W.q_norm, W.k_norm, nullptr, nullptr};
      for (int i = 0; i < 4; ++i) {
        g.sections[i].type = types[i];
        g.sections[i].out = qkvg + i * Md;
        g.sections[i].norm_weight = nw[i];
      }
      if (int rc = echo_gemm_bf16(&g, st)) return rc;
    }
    {  // joint attention over self | prefix | text | speaker keys, gated (model.py:234-264)
      echo_attn_desc a;
      memset(&a, 0, sizeof(a));
      a.q = qkvg; a.q_row_stride = d; a.q_batch_stride = (int64_t)D->seq * d;
      a.gate = qkvg + 3 * Md;
      a.out = D->ao;
      a.rows = D->rows; a.seq = D->seq; a.heads = D->heads; a.head_dim = D->head_dim;
      a.softmax_scale = (float)pow((double)D->head_dim, -0.5);
      a.impl = D->attn_impl;
      a.workspace = D->attn_workspace; a.workspace_bytes = D->attn_workspace_bytes;
      int ns = 0;
      echo_attn_segment* s = &a.seg[ns++];
      s->k = qkvg + Md; s->v = qkvg + 2 * Md;
      s->row_stride = d; s->batch_stride = (int64_t)D->seq * d; s->len = D->seq;
      if (D->prefix.len > 0) kv_segment(&a.seg[ns++], D->prefix, l);
      kv_segment(&a.seg[ns++], D->text, l);
      kv_segment(&a.seg[ns++], D->speaker, l);
      a.n_seg = ns;
      if (int rc = echo_attention(&a, st)) return rc;
    }
    if (int rc = residual_projection(D, sh, D->ao, d, W.w_o, l, 0, st)) return rc;
    if (int rc = adaln_norm(D, sh, l, 1, D->k_splits > 1, st)) return rc;
    {  // SwiGLU up-projection (model.py:308)
      echo_gemm_desc g;
      gemm_base(D, &g, D->xn, d, M, d, W.w13, 2 * D->ffn, ECHO_EPI_SWIGLU);
      g.out_bf16 = D->hmid; g.ld_out_bf16 = D->ffn;
      if (int rc = echo_gemm_bf16(&g, st)) return rc;
    }
    if (int rc = residual_projection(D, sh, D->hmid, D->ffn, W.w2, l, 1, st)) return rc;
    pending = D->k_splits > 1;
  }
  if (D->pending_out) *D->pending_out = pending ? 1 : 0;

  if (D->tail) {  // out_norm + out_proj + CFG combine + rescale + Euler (model.py:601-604, inference.py:514-558)
    echo_tail_desc t = *D->tail;
    ECHO_CHECK_ARG(t.branches > 0 && (int64_t)t.positions * t.branches == M,
                   "echo_dit_step: tail positions * branches (%d * %d) != rows * seq (%d)", t.positions, t.branches, M);
    ECHO_CHECK_ARG(t.d == d && t.latent == D->latent, "echo_dit_step: tail d / latent mismatch");
    t.h = D->x; t.ldh = d;
    t.delta = pending ? D->delta : nullptr;
    t.n_delta = pending ? D->k_splits : 0;
    t.delta_stride = Md;
    if (int rc = echo_cfg_euler_tail(&t, st)) return rc;
  }
  return ECHO_OK;
}

extern "C" int echo_dit_sample(const echo_dit_step_desc* cfg_step, const echo_dit_step_desc* plain_step, int n_steps,
                               const uint8_t* has_cfg, int64_t mod_step_stride, int64_t kv_scale_step_stride,
                               int64_t sched_step_stride, echo_stream_t stream) {
  ECHO_CHECK_ARG(n_steps >= 0 && (n_steps == 0 || has_cfg), "echo_dit_sample: null has_cfg");
  for (int i = 0; i < n_steps; ++i) {
    const echo_dit_step_desc* T = has_cfg[i] ? cfg_step : plain_step;
    ECHO_CHECK_ARG(T && T->tail && T->tail->sched && T->tail->x_t,
                   "echo_dit_sample: step %d needs a %s descriptor with a tail (sched, x_t)", i,
                   has_cfg[i] ? "three-branch" : "single-branch");
    echo_dit_step_desc s = *T;
    echo_tail_desc t = *T->tail;
    s.mod = T->mod + (int64_t)i * mod_step_stride;
    if (s.speaker.kv_scale) s.speaker.kv_scale += (int64_t)i * kv_scale_step_stride;
    t.sched = T->tail->sched + (int64_t)i * sched_step_stride;
    s.tail = &t;
    if (int rc = echo_dit_step(&s, stream)) return rc;
  }
  return ECHO_OK;
}
