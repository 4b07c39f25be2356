// Conditioning encoders as one native call (host code only): TextEncoder / SpeakerEncoder / latent-prefix encoder stack
// followed by the per-decoder-layer K/V projections.  Reference: EchoDiT.get_kv_cache_text / _speaker / _latent
// model.py:606-636, EncoderTransformerBlock model.py:311-339, SelfAttention model.py:106-161,
// JointAttention.get_kv_cache_* model.py:270-293.  `past` > 0 is the incremental form (causal encoders only).
#include <string.h>
#include "common.cuh"
#include "../../include/echo_b200.h"

namespace {

inline int64_t align256(int64_t v) { return (v + 255) / 256 * 256; }

struct Bump {
  uint8_t* base;
  int64_t off = 0;
  template <typename T>
  T* take(int64_t elems) {
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += align256(elems * (int64_t)sizeof(T));
    return p;
  }
};

void gemm_init(const echo_encode_kv_desc* D, echo_gemm_desc* g, const void* a, int64_t lda, int rows, int k, const void* w,
               int n, int epilogue) {
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

int check(const echo_encode_kv_desc* D) {
  ECHO_CHECK_ARG(D != nullptr, "encode_kv: null descriptor");
  ECHO_CHECK_ARG(D->n_layers > 0 && D->layers && D->d > 0 && D->heads > 0 && D->d == D->heads * 128 && D->ffn % 64 == 0 && D->ffn > 0,
                 "encode_kv: bad encoder shape (head_dim must be 128)");
  ECHO_CHECK_ARG(D->rows > 0 && D->seq > 0 && D->past >= 0, "encode_kv: bad rows / seq / past");
  ECHO_CHECK_ARG(D->past == 0 || (D->causal && D->enc_cache), "encode_kv: incremental encoding needs a causal encoder and enc_cache");
  ECHO_CHECK_ARG(!D->enc_cache || D->enc_cap >= D->past + D->seq, "encode_kv: enc_cap %d < past + seq", D->enc_cap);
  ECHO_CHECK_ARG(D->dec_layers > 0 && D->w_kv && D->k_norm && D->kv_out && D->dec_d == D->dec_heads * 128, "encode_kv: bad projection arguments");
  ECHO_CHECK_ARG(D->out_cap >= D->past + D->seq, "encode_kv: out_cap %d < past + seq", D->out_cap);
  ECHO_CHECK_ARG(D->rope_cos && D->rope_sin && D->pos_stride >= 1 && D->rope_heads >= 0 && D->rope_heads <= D->dec_heads, "encode_kv: bad rope arguments");
  const int pmax = D->past + D->seq - 1;
  ECHO_CHECK_ARG(pmax < D->rope_len && (D->rope_heads == 0 || pmax * D->pos_stride < D->rope_len), "encode_kv: rope table too short");
  return ECHO_OK;
}

int scatter_rows(bf16* dst, int64_t dst_pitch_elems, const bf16* src, int64_t row_elems, int rows, cudaStream_t st) {
  ECHO_CUDA(cudaMemcpy2DAsync(dst, (size_t)dst_pitch_elems * 2, src, (size_t)row_elems * 2, (size_t)row_elems * 2, (size_t)rows,
                              cudaMemcpyDeviceToDevice, st));
  return ECHO_OK;
}

// dry: workspace == nullptr, only sizes.  All prompt rows go through every launch together.  When the new positions of
// consecutive rows are not adjacent in a cache (incremental calls: row b's new keys live at [b][past .. past+seq) of a
// [rows][cap] buffer) the GEMM writes them densely to scratch and one strided copy per tensor scatters them.
int walk(const echo_encode_kv_desc* D, void* workspace, cudaStream_t st, int64_t* need) {
  const bool dry = workspace == nullptr;
  const int d = D->d, H = D->heads, L = D->seq, P0 = D->past, R = D->rows;
  const int64_t cap = D->enc_cap, ocap = D->out_cap;
  const bool enc_direct = D->enc_cache && cap == L && P0 == 0;
  const bool out_direct = ocap == L && P0 == 0;
  const int M = R * L;
  Bump ws{static_cast<uint8_t*>(workspace)};
  bf16* xn = ws.take<bf16>((int64_t)M * d);
  bf16* q = ws.take<bf16>((int64_t)M * d);
  bf16* gate = ws.take<bf16>((int64_t)M * d);
  bf16* ao = ws.take<bf16>((int64_t)M * d);
  bf16* hmid = ws.take<bf16>((int64_t)M * D->ffn);
  bf16* kscr = enc_direct ? nullptr : ws.take<bf16>((int64_t)M * d);
  bf16* vscr = enc_direct ? nullptr : ws.take<bf16>((int64_t)M * d);
  bf16* koscr = out_direct ? nullptr : ws.take<bf16>((int64_t)M * D->dec_d);
  bf16* voscr = out_direct ? nullptr : ws.take<bf16>((int64_t)M * D->dec_d);
  if (need) *need = ws.off + 256;
  if (dry) return ECHO_OK;
  bf16* cache = static_cast<bf16*>(D->enc_cache);
  bf16* out = static_cast<bf16*>(D->kv_out);
  float* x = D->x;
  for (int l = 0; l < D->n_layers; ++l) {
    const echo_enc_layer& lw = D->layers[l];
    // this layer's encoder K / V: [rows][cap][d] inside the cache, or dense scratch [rows][seq][d]
    bf16* kc = cache ? cache + ((int64_t)l * 2 + 0) * R * cap * d : kscr;
    bf16* vc = cache ? cache + ((int64_t)l * 2 + 1) * R * cap * d : vscr;
    const int64_t kcap = cache ? cap : L;
    bf16* ksec = enc_direct ? kc : kscr;
    bf16* vsec = enc_direct ? vc : vscr;
    if (int rc = echo_rmsnorm_mod(x, d, M, d, D->eps, lw.attn_norm, nullptr, 0, 0, nullptr, 0, xn, d, st)) return rc;
    {
      echo_gemm_desc g;
      gemm_init(D, &g, xn, d, M, d, lw.w_qkvg, 4 * d, ECHO_EPI_SECTIONS);
      g.n_sections = 4; g.section_width = d; g.head_dim = 128; g.rope_heads = H;
      g.pos_offset = P0; g.pos_stride = 1; g.rows_per_seq = L; g.eps = D->eps;
      g.rope_cos = D->rope_cos; g.rope_sin = D->rope_sin; g.ld_sec_out = d;
      g.sections[0].type = ECHO_SEC_NORM_ROPE; g.sections[0].out = q; g.sections[0].norm_weight = lw.q_norm;
      g.sections[1].type = ECHO_SEC_NORM_ROPE; g.sections[1].out = ksec; g.sections[1].norm_weight = lw.k_norm;
      g.sections[2].type = ECHO_SEC_PLAIN; g.sections[2].out = vsec;
      g.sections[3].type = ECHO_SEC_SIGMOID; g.sections[3].out = gate;
      if (int rc = echo_gemm_bf16(&g, st)) return rc;
    }
    if (cache && !enc_direct) {
      if (int rc = scatter_rows(kc + (int64_t)P0 * d, cap * d, kscr, (int64_t)L * d, R, st)) return rc;
      if (int rc = scatter_rows(vc + (int64_t)P0 * d, cap * d, vscr, (int64_t)L * d, R, st)) return rc;
    }
    {
      echo_attn_desc a;
      memset(&a, 0, sizeof(a));
      a.q = q; a.q_row_stride = d; a.q_batch_stride = (int64_t)L * d;
      a.gate = gate; a.out = ao;
      a.rows = R; a.seq = L; a.heads = H; a.head_dim = 128;
      a.n_seg = 1; a.impl = 0;
      a.softmax_scale = 1.0f / sqrtf(128.0f);
      echo_attn_segment& s = a.seg[0];
      s.k = kc; s.v = vc; s.row_stride = d; s.batch_stride = kcap * d;
      s.len = P0 + L; s.causal = D->causal; s.q_offset = P0;
      s.mask = D->mask; s.mask_stride = D->mask_stride; s.extent = D->extent;
      if (int rc = echo_attention(&a, st)) return rc;
    }
    {
      echo_gemm_desc g;
      gemm_init(D, &g, ao, d, M, d, lw.w_o, d, ECHO_EPI_GENERIC);
      g.resid = x; g.ld_resid = d; g.out_f32 = x; g.ld_out_f32 = d;
      if (int rc = echo_gemm_bf16(&g, st)) return rc;
    }
    if (int rc = echo_rmsnorm_mod(x, d, M, d, D->eps, lw.mlp_norm, nullptr, 0, 0, nullptr, 0, xn, d, st)) return rc;
    {
      echo_gemm_desc g;
      gemm_init(D, &g, xn, d, M, d, lw.w13, 2 * D->ffn, ECHO_EPI_SWIGLU);
      g.out_bf16 = hmid; g.ld_out_bf16 = D->ffn;
      if (int rc = echo_gemm_bf16(&g, st)) return rc;
    }
    {
      echo_gemm_desc g;
      gemm_init(D, &g, hmid, D->ffn, M, D->ffn, lw.w2, d, ECHO_EPI_GENERIC);
      g.resid = x; g.ld_resid = d; g.out_f32 = x; g.ld_out_f32 = d;
      if (int rc = echo_gemm_bf16(&g, st)) return rc;
    }
  }
  // text_norm / speaker_norm / latent_norm, then the 2 * dec_layers projections (K normalised per head, prefix K rotated)
  if (int rc = echo_rmsnorm_mod(x, d, M, d, D->eps, D->out_norm, nullptr, 0, 0, nullptr, 0, xn, d, st)) return rc;
  const int dd = D->dec_d;
  for (int l = 0; l < D->dec_layers; ++l) {
    bf16* ko = out + ((int64_t)l * 2 + 0) * R * ocap * dd;
    bf16* vo = out + ((int64_t)l * 2 + 1) * R * ocap * dd;
    echo_gemm_desc g;
    gemm_init(D, &g, xn, d, M, d, D->w_kv[l], 2 * dd, ECHO_EPI_SECTIONS);
    g.n_sections = 2; g.section_width = dd; g.head_dim = 128; g.rope_heads = D->rope_heads;
    g.pos_offset = P0 * D->pos_stride; g.pos_stride = D->pos_stride; g.rows_per_seq = L; g.eps = D->eps;
    g.rope_cos = D->rope_cos; g.rope_sin = D->rope_sin; g.ld_sec_out = dd;
    g.sections[0].type = ECHO_SEC_NORM_ROPE; g.sections[0].out = out_direct ? ko : koscr; g.sections[0].norm_weight = D->k_norm[l];
    g.sections[1].type = ECHO_SEC_PLAIN; g.sections[1].out = out_direct ? vo : voscr;
    if (int rc = echo_gemm_bf16(&g, st)) return rc;
    if (!out_direct) {
      if (int rc = scatter_rows(ko + (int64_t)P0 * dd, ocap * dd, koscr, (int64_t)L * dd, R, st)) return rc;
      if (int rc = scatter_rows(vo + (int64_t)P0 * dd, ocap * dd, voscr, (int64_t)L * dd, R, st)) return rc;
    }
  }
  return ECHO_OK;
}

}  // namespace

extern "C" int64_t echo_encode_kv_workspace_bytes(const echo_encode_kv_desc* D) {
  if (check(D) != ECHO_OK) return -1;
  int64_t need = 0;
  walk(D, nullptr, nullptr, &need);
  return need;
}

extern "C" int echo_encode_kv(const echo_encode_kv_desc* D, void* workspace, int64_t workspace_bytes, echo_stream_t stream) {
  if (int rc = check(D)) return rc;
  ECHO_CHECK_ARG(D->x && workspace && ((uintptr_t)workspace % 256) == 0, "encode_kv: null / misaligned x or workspace");
  int64_t need = 0;
  walk(D, nullptr, nullptr, &need);
  ECHO_CHECK_ARG(workspace_bytes >= need, "encode_kv: workspace too small (%lld bytes, need %lld)", (long long)workspace_bytes, (long long)need);
  return walk(D, workspace, (cudaStream_t)stream, nullptr);
}
