// Fused GEMM epilogues shared by the tcgen05 kernel (accumulator read from TMEM) and the legacy
// mma.sync kernel (accumulator staged in shared memory).  One thread owns one output row and walks
// it in 32-column chunks through a Loader:  loader.load32(col_in_tile, float v[32]).
#pragma once
#include "common.cuh"
#include "../../include/echo_b200.h"

struct GemmShape {
  int k_blocks_per_tap, taps, k_tap;
  int tap_shift[8];
  int rows, n;
  int a_ctx;                 // rows of left context in memory before row 0 (already folded into a and tap_shift)
  long long lda, a_batch_stride, ldb;
  const bf16* a;
  const bf16* b;
};

struct EpiSection {
  int type;
  bf16* out;
  const float* norm_w;
};

struct EpiParams {
  int kind;
  int rows, n;               // rows per batch, total N
  // generic
  int act;
  const float* bias;
  const float* gate; long long gate_ld; int rows_per_gate;
  const float* resid; long long ld_resid;
  int red_resid;             // resid aliases out_f32 and nothing else consumes the sum: the pair kernel issues
                             // x += delta as a vector reduction (no residual load on the epilogue's critical path)
  float* out_f32; long long ld_out_f32;
  bf16* out_bf16; long long ld_out_bf16;
  bf16* out2_bf16; long long ld_out2_bf16;
  const float* snake_alpha;
  // sections
  int n_sections, section_width, head_dim, rope_heads;
  int pos_offset, pos_stride, rows_per_seq;
  float eps;
  const float* rope_cos; const float* rope_sin;
  long long ld_sec_out;
  EpiSection sec[4];
};

__device__ __forceinline__ void store_bf16x8(bf16* p, const float* v) {
  uint4 u;
  u.x = pack_bf16x2(v[0], v[1]); u.y = pack_bf16x2(v[2], v[3]);
  u.z = pack_bf16x2(v[4], v[5]); u.w = pack_bf16x2(v[6], v[7]);
  *reinterpret_cast<uint4*>(p) = u;
}

// ---- GENERIC ------------------------------------------------------------------------------
template <class Loader>
__device__ __forceinline__ void epi_generic(const int BN, const EpiParams& P, Loader& L, bool row_ok, long long m_glob,
                                            int n0) {
  const float* gate_row = P.gate ? P.gate + (m_glob / P.rows_per_gate) * P.gate_ld : nullptr;
#pragma unroll 1
  for (int c = 0; c < BN; c += 32) {
    float v[32];
    L.load32(c, v);
    if (!row_ok) continue;
    const int nb = n0 + c;
    if (nb >= P.n) continue;
#pragma unroll
    for (int g = 0; g < 32; g += 8) {
      const int n = nb + g;
      if (n >= P.n) break;
      float* w = v + g;
      if (P.bias) {
#pragma unroll
        for (int j = 0; j < 8; ++j) w[j] += __ldg(P.bias + n + j);
      }
      if (P.act == ECHO_ACT_GELU) {
#pragma unroll
        for (int j = 0; j < 8; ++j) w[j] = geluf_(w[j]);
      } else if (P.act == ECHO_ACT_SILU) {
#pragma unroll
        for (int j = 0; j < 8; ++j) w[j] = siluf_(w[j]);
      }
      if (gate_row) {
#pragma unroll
        for (int j = 0; j < 8; ++j) w[j] *= __ldg(gate_row + n + j);
      }
      if (P.resid) {
        const float4* r = reinterpret_cast<const float4*>(P.resid + m_glob * P.ld_resid + n);
        float4 r0 = r[0], r1 = r[1];
        w[0] += r0.x; w[1] += r0.y; w[2] += r0.z; w[3] += r0.w;
        w[4] += r1.x; w[5] += r1.y; w[6] += r1.z; w[7] += r1.w;
      }
      if (P.out_f32) {
        float4* o = reinterpret_cast<float4*>(P.out_f32 + m_glob * P.ld_out_f32 + n);
        o[0] = make_float4(w[0], w[1], w[2], w[3]);
        o[1] = make_float4(w[4], w[5], w[6], w[7]);
      }
      if (P.out_bf16) store_bf16x8(P.out_bf16 + m_glob * P.ld_out_bf16 + n, w);
      if (P.out2_bf16) {
        float s[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) s[j] = snakef_(w[j], __ldg(P.snake_alpha + n + j));
        store_bf16x8(P.out2_bf16 + m_glob * P.ld_out2_bf16 + n, s);
      }
    }
  }
}

// ---- SECTIONS (q/k/v/gate style) ------------------------------------------------------------
template <class Loader>
__device__ __forceinline__ void epi_sections(const int BN, const EpiParams& P, Loader& L, bool row_ok, long long m_glob,
                                             int m_in_batch, int n0) {
  const int sec_idx = n0 / P.section_width;
  const EpiSection S = P.sec[sec_idx];
  const int sec_col0 = n0 - sec_idx * P.section_width;   // column of the tile inside its section
  const int hd = P.head_dim;
  const int pos = P.pos_offset + (m_in_batch % P.rows_per_seq) * P.pos_stride;
  bf16* out_row = S.out + m_glob * P.ld_sec_out + sec_col0;
#pragma unroll 1
  for (int h0 = 0; h0 < BN; h0 += hd) {             // one head at a time
    const int head = (sec_col0 + h0) / hd;
    const bool in_range = (n0 + h0) < P.n;
    float rstd = 1.f;
    if (S.type == ECHO_SEC_NORM_ROPE) {
      float ssq = 0.f;
#pragma unroll 1
      for (int c = 0; c < hd; c += 32) {
        float v[32];
        L.load32(h0 + c, v);
#pragma unroll
        for (int j = 0; j < 32; ++j) ssq += v[j] * v[j];
      }
      rstd = rsqrtf(ssq / (float)hd + P.eps);
    }
    const bool do_rope = (S.type == ECHO_SEC_NORM_ROPE || S.type == ECHO_SEC_ROPE) && head < P.rope_heads;
#pragma unroll 1
    for (int c = 0; c < hd; c += 32) {
      float v[32];
      L.load32(h0 + c, v);
      if (!row_ok || !in_range) continue;
      if (S.type == ECHO_SEC_NORM_ROPE) {
        const float* w = S.norm_w + sec_col0 + h0 + c;
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = v[j] * rstd * __ldg(w + j);
      } else if (S.type == ECHO_SEC_SIGMOID) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = sigmoidf_(v[j]);
      }
      if (do_rope) {
        const float* cs = P.rope_cos + (long long)pos * (hd / 2) + c / 2;
        const float* sn = P.rope_sin + (long long)pos * (hd / 2) + c / 2;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float co = __ldg(cs + j), si = __ldg(sn + j);
          const float a = v[2 * j], b = v[2 * j + 1];
          v[2 * j] = a * co - b * si;
          v[2 * j + 1] = a * si + b * co;
        }
      }
#pragma unroll
      for (int g = 0; g < 32; g += 8) store_bf16x8(out_row + h0 + c + g, v + g);
    }
  }
}

// ---- SWIGLU ---------------------------------------------------------------------------------
template <class Loader>
__device__ __forceinline__ void epi_swiglu(const int BN, const EpiParams& P, Loader& L, bool row_ok, long long m_glob, int n0) {
#pragma unroll 1
  for (int g0 = 0; g0 < BN; g0 += 128) {
    const int out_col0 = (n0 + g0) / 2;
#pragma unroll 1
    for (int c = 0; c < 64; c += 32) {
      float a[32], b[32];
      L.load32(g0 + c, a);
      L.load32(g0 + 64 + c, b);
      if (!row_ok || (n0 + g0) >= P.n) continue;
#pragma unroll
      for (int j = 0; j < 32; ++j) a[j] = siluf_(a[j]) * b[j];
      bf16* o = P.out_bf16 + m_glob * P.ld_out_bf16 + out_col0 + c;
#pragma unroll
      for (int g = 0; g < 32; g += 8) store_bf16x8(o + g, a + g);
    }
  }
}

// BN (tile width in columns, multiple of 32) is a run-time value: the column loops are not unrolled.
template <int EPI, class Loader>
__device__ __forceinline__ void run_epilogue(const int BN, const EpiParams& P, Loader& L, int batch_idx, int m_in_batch,
                                             int n0) {
  const bool row_ok = m_in_batch < P.rows;
  const long long m_glob = (long long)batch_idx * P.rows + m_in_batch;
  if (EPI == ECHO_EPI_GENERIC) epi_generic(BN, P, L, row_ok, m_glob, n0);
  else if (EPI == ECHO_EPI_SECTIONS) epi_sections(BN, P, L, row_ok, m_glob, m_in_batch, n0);
  else epi_swiglu(BN, P, L, row_ok, m_glob, n0);
}
