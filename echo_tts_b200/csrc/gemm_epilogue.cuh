// Kernel-side parameter blocks of the fused GEMM epilogues (gemm_epilogue_t.cuh), filled from echo_gemm_desc by gemm.cu.
#pragma once
#include "common.cuh"
#include "../../include/echo_b200.h"

struct GemmShape {
  int k_blocks_per_tap, taps, k_tap;
  int tap_shift[8];
  int rows, n;
  int a_ctx;                 // rows of left context in memory before row 0 (already folded into a and tap_shift)
  int b_static;              // echo_gemm_desc.b_static: B may be requested before the grid dependency resolves
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
