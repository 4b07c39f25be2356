// Coalesced, vectorised, latency-tolerant fused epilogues for the CTA-pair GEMM.
//
// The accumulator arrives from TMEM "row per thread" (tcgen05.ld 32x32b: lane r holds 32 consecutive columns of
// row r).  Math that needs a whole row (per-head sum of squares, sigmoid, SwiGLU pairing) is done in that layout.
// Everything that touches global memory (bias / gate / residual / RoPE table / outputs) is done after a 32x32
// transpose through an XOR-swizzled per-warp shared-memory scratch in a "4 columns x 8 row-groups per lane"
// layout: lane l owns columns 4*(l&7)..+3 of rows (l>>3) + 4*i, i = 0..7, so that each warp-wide 128-bit access
// covers four complete 128-byte row segments (fully used sectors, 8 instructions per 32x32 chunk).
//
// While the GEMM main loop saturates the L2 fabric a global load takes 2-3 k cycles, so no load may sit on the
// per-chunk critical path: each of the EIGHT epilogue warps (two per TMEM lane quarter, alternating 32-column
// chunks) issues the auxiliary loads of its NEXT chunk (residual rows, RoPE rows, per-column vectors) before it
// starts on the current one.  The residual may alias the output, so the compiler cannot do this hoisting itself.
#pragma once
#include "gemm_epilogue.cuh"

constexpr int EPI_SC_FLOATS = 32 * 32;
// Kernel-internal epilogue id: GENERIC without activation / bf16 / Snake outputs and with the residual either absent or
// updated in place (the DiT's two residual projections and their split-K form): leaves through TMA (epi_t_lean_tma).
constexpr int EPI_GENERIC_LEAN = 3;
#ifdef ECHO_DEBUG
// timing experiments of the debug library only (RESULTS ARE GARBAGE when g_epi_dbg != 0): bit 0 = no bulk store issued,
// bit 1 = no gate / bias shuffles, bit 2 = no staging stores; g_epi_trace: [cta][32] clock stamps of epilogue warp 0
__device__ int g_epi_dbg = 0;
__device__ long long* g_epi_trace = nullptr;
#define EPI_STAMP(k) do { if (etr && (k) < 32) etr[(k)] = clock64(); } while (0)
#else
#define EPI_STAMP(k) do { } while (0)
#endif               // per-warp scratch: 32 rows x 32 fp32, 16-byte chunks swizzled

// The scratch is addressed in the shared state space explicitly: through a generic pointer the compiler emits
// generic LD/ST (address translation, long-scoreboard latency) instead of LDS/STS.
// -DECHO_DEBUG_EPI_NOSMEM (timing experiments only, results garbage): the transpose leaves shared memory alone, which
// shows how much of the main loop's shared-memory bandwidth the overlapped epilogues take.
__device__ __forceinline__ void sc_store_row(float* sc, int lane, const float* v) {
#ifdef ECHO_DEBUG_EPI_NOSMEM
  return;
#endif
  const uint32_t row = smem_u32(sc) + (uint32_t)lane * 128u;
#pragma unroll
  for (int k = 0; k < 8; ++k)
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(row + (uint32_t)((k ^ (lane & 7)) << 4)), "f"(v[4 * k]),
                 "f"(v[4 * k + 1]), "f"(v[4 * k + 2]), "f"(v[4 * k + 3])
                 : "memory");
  __syncwarp();
}
__device__ __forceinline__ float4 sc_load4(const float* sc, int row, int cg) {
  float4 x;
#ifdef ECHO_DEBUG_EPI_NOSMEM
  x = make_float4((float)row, (float)cg, 0.f, 0.f);
  return x;
#endif
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(x.x), "=f"(x.y), "=f"(x.z), "=f"(x.w)
               : "r"(smem_u32(sc) + (uint32_t)(row * 128 + ((cg ^ (row & 7)) << 4)))
               : "memory");
  return x;
}
// x[0..3] += v as ONE fire-and-forget vector reduction executed in the L2 (sm_90+): the in-place residual update
// without reading the residual in the epilogue (its rows are usually back in DRAM by then: ~3 k clocks per chunk on
// the exposed last epilogue of every launch)
__device__ __forceinline__ void red_add_f32x4(float* p, float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void store_bf16x4(bf16* p, float4 x) {
  *reinterpret_cast<uint2*>(p) = make_uint2(pack_bf16x2(x.x, x.y), pack_bf16x2(x.z, x.w));
}

// Rows handled by this warp: m_in_batch = mb0 + r, r = 0..31 (lane r owns row r in TMEM layout).
struct EpiRows {
  int batch_idx, mb0, n0;
  int nrows;            // valid rows among the 32 (0..32), warp-uniform
  int half;             // 0/1: which of the two warps of this lane quarter (chunk parity it owns)
  int split;            // split-K unit index (bias only in unit 0)
  long long out_off;    // split * split_stride: offset of this unit's partial output
  long long mg0;        // global row index of row 0
};

// ---- GENERIC ------------------------------------------------------------------------------
// Only the residual rows (one 128-byte segment of 32 different rows: L2 latency) are fetched a chunk ahead; the
// per-column vectors (bias / gate / snake alpha: a few KB shared by every row, L1 hits after the first touch) are
// requested at the top of process(), under the TMEM load and the transpose.
struct GenAux {
  float4 rv[8];
};

struct GenCtx {
  const EpiParams& P;
  const EpiRows& R;
  int cg, rsub, rpg, rem0;
  long long g0;
  bool gate_fast;

  __device__ __forceinline__ void load(int c, GenAux& A) const {
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    const int n = R.n0 + c + 4 * cg;
    const bool nok = n < P.n;                            // n % 8 == 0: a 4-column group is all in or all out
    const float* rp = (P.resid && !P.red_resid) ? P.resid + R.mg0 * P.ld_resid + n : nullptr;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = 4 * i + rsub;
      A.rv[i] = (rp && nok && r < R.nrows) ? *reinterpret_cast<const float4*>(rp + (long long)r * P.ld_resid) : zero4;
    }
  }

  template <bool LEAN, class Loader>
  __device__ __forceinline__ void process(int c, const GenAux& A, Loader& L, float* sc, int lane) const {
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f), one4 = make_float4(1.f, 1.f, 1.f, 1.f);
    const int n = R.n0 + c + 4 * cg;
    const bool nok = n < P.n;
    const bool two_gates = rem0 + 31 >= rpg;             // warp-uniform: the 32 rows straddle a gate-row boundary
    const float4 bias = (P.bias && nok && R.split == 0) ? __ldg(reinterpret_cast<const float4*>(P.bias + n)) : zero4;
    float4 glo = one4, ghi = one4;
    if (P.gate && nok && gate_fast) {
      glo = __ldg(reinterpret_cast<const float4*>(P.gate + g0 * P.gate_ld + n));
      if (two_gates) ghi = __ldg(reinterpret_cast<const float4*>(P.gate + (g0 + 1) * P.gate_ld + n));
    }
    float v[32];
    L.load32(c, v);
    sc_store_row(sc, lane, v);
    float4 x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = sc_load4(sc, 4 * i + rsub, cg);
    if ((LEAN || P.act == ECHO_ACT_NONE) && (!P.gate || (gate_fast && !two_gates))) {
      // the DiT projections: x = (acc + bias) * g + resid as one FMA per element
      const float4 bg = make_float4(bias.x * glo.x, bias.y * glo.y, bias.z * glo.z, bias.w * glo.w);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        x[i].x = fmaf(x[i].x, glo.x, bg.x + A.rv[i].x); x[i].y = fmaf(x[i].y, glo.y, bg.y + A.rv[i].y);
        x[i].z = fmaf(x[i].z, glo.z, bg.z + A.rv[i].z); x[i].w = fmaf(x[i].w, glo.w, bg.w + A.rv[i].w);
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) { x[i].x += bias.x; x[i].y += bias.y; x[i].z += bias.z; x[i].w += bias.w; }
      if (!LEAN && P.act == ECHO_ACT_GELU) {
#pragma unroll
        for (int i = 0; i < 8; ++i) { x[i].x = geluf_(x[i].x); x[i].y = geluf_(x[i].y); x[i].z = geluf_(x[i].z); x[i].w = geluf_(x[i].w); }
      } else if (!LEAN && P.act == ECHO_ACT_SILU) {
#pragma unroll
        for (int i = 0; i < 8; ++i) { x[i].x = siluf_(x[i].x); x[i].y = siluf_(x[i].y); x[i].z = siluf_(x[i].z); x[i].w = siluf_(x[i].w); }
      }
      if (P.gate) {
        if (gate_fast) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 g = (rem0 + 4 * i + rsub >= rpg) ? ghi : glo;
            x[i].x *= g.x; x[i].y *= g.y; x[i].z *= g.z; x[i].w *= g.w;
          }
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int r = 4 * i + rsub;
            if (nok && r < R.nrows) {
              const float4 g = __ldg(reinterpret_cast<const float4*>(P.gate + ((R.mg0 + r) / rpg) * P.gate_ld + n));
              x[i].x *= g.x; x[i].y *= g.y; x[i].z *= g.z; x[i].w *= g.w;
            }
          }
        }
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) { x[i].x += A.rv[i].x; x[i].y += A.rv[i].y; x[i].z += A.rv[i].z; x[i].w += A.rv[i].w; }
    }
    if (P.red_resid) {       // x[i] holds the update only (A.rv is zero): out_f32 += x
      float* o = P.out_f32 + R.mg0 * P.ld_out_f32 + n;
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (nok && 4 * i + rsub < R.nrows) red_add_f32x4(o + (long long)(4 * i + rsub) * P.ld_out_f32, x[i]);
    } else if (P.out_f32) {
      float* o = P.out_f32 + R.out_off + R.mg0 * P.ld_out_f32 + n;
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (nok && 4 * i + rsub < R.nrows) *reinterpret_cast<float4*>(o + (long long)(4 * i + rsub) * P.ld_out_f32) = x[i];
    }
    if (!LEAN && P.out_bf16) {
      bf16* o = P.out_bf16 + R.mg0 * P.ld_out_bf16 + n;
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (nok && 4 * i + rsub < R.nrows) store_bf16x4(o + (long long)(4 * i + rsub) * P.ld_out_bf16, x[i]);
    }
    if (!LEAN && P.out2_bf16) {
      bf16* o = P.out2_bf16 + R.mg0 * P.ld_out2_bf16 + n;
      const float4 alpha = nok ? __ldg(reinterpret_cast<const float4*>(P.snake_alpha + n)) : one4;
      const float4 ia = make_float4(__fdividef(1.f, alpha.x + 1e-9f), __fdividef(1.f, alpha.y + 1e-9f),
                                    __fdividef(1.f, alpha.z + 1e-9f), __fdividef(1.f, alpha.w + 1e-9f));
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float4 s;
        s.x = __sinf(alpha.x * x[i].x); s.y = __sinf(alpha.y * x[i].y);
        s.z = __sinf(alpha.z * x[i].z); s.w = __sinf(alpha.w * x[i].w);
        s.x = fmaf(s.x * s.x, ia.x, x[i].x); s.y = fmaf(s.y * s.y, ia.y, x[i].y);
        s.z = fmaf(s.z * s.z, ia.z, x[i].z); s.w = fmaf(s.w * s.w, ia.w, x[i].w);
        if (nok && 4 * i + rsub < R.nrows) store_bf16x4(o + (long long)(4 * i + rsub) * P.ld_out2_bf16, s);
      }
    }
    __syncwarp();
  }
};

__device__ __forceinline__ GenCtx gen_ctx(const EpiParams& P, const EpiRows& R, int lane) {
  // gate row of local row r is g0 + ((rem0 + r) >= rows_per_gate) when rows_per_gate >= 32
  const int rpg = P.rows_per_gate;
  const unsigned g0 = (unsigned)R.mg0 / (unsigned)rpg;    // rows fit 32 bits (checked at launch); avoids the 64-bit divide routine
  return GenCtx{P, R, lane & 7, lane >> 3, rpg, (int)((unsigned)R.mg0 - g0 * (unsigned)rpg), (long long)g0, rpg >= 32};
}
// The first chunk's residual rows do not depend on the accumulator: they are requested BEFORE the warp waits for the
// tile's MMAs, so their latency is off the exposed epilogue of the last tile (and of every other one).
__device__ __forceinline__ void epi_t_generic_prefetch(const int BN, const EpiParams& P, int lane, const EpiRows& R,
                                                       GenAux& A) {
  if (R.nrows <= 0) return;
  const GenCtx C = gen_ctx(P, R, lane);
  if (R.half * 32 < min(BN, P.n - R.n0)) C.load(R.half * 32, A);
}
template <bool LEAN, class Loader>
__device__ __forceinline__ void epi_t_generic(const int BN, const EpiParams& P, Loader& L, float* sc, int lane,
                                              const EpiRows& R, GenAux& A) {
  if (R.nrows <= 0) return;                               // warp-uniform; nothing was promised to TMEM readers
  const GenCtx C = gen_ctx(P, R, lane);
  const int cend = min(BN, P.n - R.n0);                   // columns of this tile that exist (multiple of 8)
  // this warp's chunks: half*32, +64, ...  ONE rolled call site of process(): the ~300-instruction body then stays in
  // the instruction cache from the second chunk on (the unrolled two-site version spent 20 % of the exposed epilogue on
  // instruction fetch).  A arrives pre-loaded with the first chunk's residual rows (epi_t_generic_prefetch); chunk
  // k+1's rows are requested into B before chunk k is processed.
  int c = R.half * 32;
  if (P.red_resid) {                                     // nothing to fetch ahead: x += update goes out as reductions
#pragma unroll
    for (int i = 0; i < 8; ++i) A.rv[i] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
    for (; c < cend; c += 64) C.template process<LEAN>(c, A, L, sc, lane);
    return;
  }
#pragma unroll 1
  for (; c < cend; c += 64) {
    GenAux B;
    const bool has_next = c + 64 < cend;
    if (has_next) C.load(c + 64, B);
    C.template process<LEAN>(c, A, L, sc, lane);
    if (has_next) A = B;
  }
}

// ---- SECTIONS (q/k/v/gate style) ------------------------------------------------------------
struct SecAux {
  float4 w;
  float2 co[8], si[8];
};

// norm weight + RoPE rows of one 32-column chunk (columns h0 + c .. of the tile) for this lane's 4 columns x 8 rows
// Per-tile invariants of the sections epilogue (one division / modulo each per tile instead of per chunk: the address
// arithmetic of sec_load_aux used to take ~1.3 k clocks per head).
struct SecTile {
  int sec_idx, sec_col0;       // section of the tile and the tile's first column inside it
  int seq0;                    // sequence index of the warp's local row 0
};
__device__ __forceinline__ SecTile sec_tile(const EpiParams& P, const EpiRows& R) {
  SecTile t;
  t.sec_idx = R.n0 / P.section_width;
  t.sec_col0 = R.n0 - t.sec_idx * P.section_width;
  t.seq0 = R.mb0 % P.rows_per_seq;
  return t;
}
__device__ __forceinline__ void sec_load_aux(const EpiParams& P, const EpiRows& R, const SecTile& T, int lane, int h0, int c,
                                             bool do_rope, SecAux& A) {
  const int cg = lane & 7, rsub = lane >> 3;
  const EpiSection& S = P.sec[T.sec_idx];
  const int col = T.sec_col0 + h0 + c + 4 * cg;             // first of this lane's 4 columns inside the section
  A.w = S.type == ECHO_SEC_NORM_ROPE ? __ldg(reinterpret_cast<const float4*>(S.norm_w + col)) : make_float4(1.f, 1.f, 1.f, 1.f);
  if (do_rope) {                                            // lane's columns = rotary pairs p0, p0 + 1 of the head
    const int p0 = (c + 4 * cg) >> 1, hp = P.head_dim >> 1;
    int sq = T.seq0 + rsub;                                 // sequence index of row r = 4 i + rsub, wrapped by subtraction
    while (sq >= P.rows_per_seq) sq -= P.rows_per_seq;
    const int step = 4 % P.rows_per_seq;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = 4 * i + rsub;
      const int off = (P.pos_offset + sq * P.pos_stride) * hp + p0;     // < 2^31: rotary tables are a few MB
      const bool ok = r < R.nrows;
      A.co[i] = ok ? __ldg(reinterpret_cast<const float2*>(P.rope_cos + off)) : make_float2(1.f, 1.f);
      A.si[i] = ok ? __ldg(reinterpret_cast<const float2*>(P.rope_sin + off)) : make_float2(0.f, 0.f);
      sq += step;
      if (sq >= P.rows_per_seq) sq -= P.rows_per_seq;
    }
  }
}
template <class Loader>
__device__ __forceinline__ void epi_t_sections(const int BN, const EpiParams& P, Loader& L, float* sc, int lane,
                                               const EpiRows& R) {
  if (R.nrows <= 0) return;
  const int cg = lane & 7, rsub = lane >> 3;
  const SecTile ST = sec_tile(P, R);
  const EpiSection S = P.sec[ST.sec_idx];
  const int sec_col0 = ST.sec_col0;                        // column of the tile inside its section
  const int hd = P.head_dim;
  const long long ld = P.ld_sec_out;
  const bool norm = S.type == ECHO_SEC_NORM_ROPE;
  int head = sec_col0 / hd;                                // head of the tile's first column; + 1 per iteration below
#ifdef ECHO_DEBUG
  // stamps of the CTA's first two tiles (slots 0..14 and 16..30; slot 31 counts the calls), epilogue warp 0 only
  long long* etr = (g_epi_trace && (threadIdx.x >> 5) == 2 && lane == 0) ? g_epi_trace + (long long)blockIdx.x * 32 : nullptr;
  if (etr) { const long long n = etr[31]++; etr = n < 2 ? etr + 16 * n : nullptr; }
  int stamp = 0;
#define SEC_STAMP() do { if (etr && stamp < 15) etr[stamp++] = clock64(); } while (0)
#else
#define SEC_STAMP() do { } while (0)
#endif
  SEC_STAMP();

  auto process = [&](int h0, int c, bool do_rope, float rstd, const SecAux& A) {
    float v[32];
    L.load32(h0 + c, v);
    if (norm) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] *= rstd;
    } else if (S.type == ECHO_SEC_SIGMOID) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = __fdividef(1.f, 1.f + __expf(-v[j]));
    }
    sc_store_row(sc, lane, v);
    bf16* out = S.out + R.mg0 * ld + sec_col0 + h0 + c + 4 * cg;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = 4 * i + rsub;
      float4 x = sc_load4(sc, r, cg);
      x.x *= A.w.x; x.y *= A.w.y; x.z *= A.w.z; x.w *= A.w.w;
      if (do_rope) {
        float4 y;
        y.x = x.x * A.co[i].x - x.y * A.si[i].x;
        y.y = x.x * A.si[i].x + x.y * A.co[i].x;
        y.z = x.z * A.co[i].y - x.w * A.si[i].y;
        y.w = x.z * A.si[i].y + x.w * A.co[i].y;
        x = y;
      }
      if (r < R.nrows) store_bf16x4(out + (long long)r * ld, x);
    }
    __syncwarp();
  };

#pragma unroll 1
  for (int h0 = 0; h0 < BN; h0 += hd, ++head) {            // one head at a time; both warps of the quarter visit it
    if (R.n0 + h0 >= P.n) break;
    const bool do_rope = (norm || S.type == ECHO_SEC_ROPE) && head < P.rope_heads;
    // this warp's chunks of the head: c = half*32 (+64 when the head is 128 wide); their aux loads go out first
    const int c0 = R.half * 32, c1 = c0 + 64;
    SecAux A0, A1;
    sec_load_aux(P, R, ST, lane, h0, c0, do_rope, A0);
    if (c1 < hd) sec_load_aux(P, R, ST, lane, h0, c1, do_rope, A1);
    SEC_STAMP();
    float rstd = 1.f;
    if (norm) {
      // row-per-thread sum of squares over the whole head: each of the quarter's two warps sums the columns it will
      // process (half the TMEM reads of a full pass per warp) and the two exchange their partial sums through
      // the first 128 bytes of their scratch (idle between chunks), fenced by a 64-thread named barrier on each side
      float ssq = 0.f;
#pragma unroll 1
      for (int c = c0; c < hd; c += 64) {                  // (keeping a chunk in registers for process() spills)
        float t[32];
        L.load32(h0 + c, t);
#pragma unroll
        for (int j = 0; j < 32; ++j) ssq = fmaf(t[j], t[j], ssq);
      }
      const int bar_id = 4 + (int)((threadIdx.x >> 5) & 3);
      const float* other = R.half ? sc - 4 * EPI_SC_FLOATS : sc + 4 * EPI_SC_FLOATS;
      sc[lane] = ssq;
      asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
      ssq += other[lane];
      asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
      rstd = rsqrtf(ssq / (float)hd + P.eps);
    }
    SEC_STAMP();
#pragma unroll 1
    for (int c = c0; c < hd; c += 64) {                   // one rolled call site (instruction-cache footprint)
      process(h0, c, do_rope, rstd, A0);
      A0 = A1;
      SEC_STAMP();
    }
  }
}

// ---- SWIGLU ---------------------------------------------------------------------------------
template <class Loader>
__device__ __forceinline__ void epi_t_swiglu(const int BN, const EpiParams& P, Loader& L, float* sc, int lane,
                                             const EpiRows& R) {
  if (R.nrows <= 0) return;
  const int cg = lane & 7, rsub = lane >> 3;
  const long long ld = P.ld_out_bf16;
  const int c = R.half * 32;                               // this warp's 32 of the 64 (w1 | w3) column pairs per group
#pragma unroll 1
  for (int g0 = 0; g0 < BN; g0 += 128) {
    if (R.n0 + g0 >= P.n) break;
    const int out_col0 = (R.n0 + g0) / 2;
    float a[32], b[32];
    L.load32(g0 + c, a);
    L.load32(g0 + 64 + c, b);
#pragma unroll
    for (int j = 0; j < 32; ++j) a[j] = __fdividef(a[j], 1.f + __expf(-a[j])) * b[j];
    sc_store_row(sc, lane, a);
    bf16* out = P.out_bf16 + R.mg0 * ld + out_col0 + c + 4 * cg;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = 4 * i + rsub;
      if (r < R.nrows) store_bf16x4(out + (long long)r * ld, sc_load4(sc, r, cg));
    }
    __syncwarp();
  }
}

// ---- LEAN via TMA (the DiT's gated-residual projections and their split-K partials) ---------------------------------
// The LEAN cases -- x += gate * (acc + bias) in place, or out = gate * (acc + bias) (split-K partial products) -- need no
// per-lane global access at all: the chunk is scaled in the row-per-thread TMEM layout, written once to the warp's
// scratch in the SWIZZLE_128B pattern (the XOR swizzle sc_store_row already uses), and leaves as ONE bulk tensor
// operation issued by lane 0: a TMA reduce-add (x += ..., executed in the L2, one add per element: deterministic) or a TMA
// store.  Rows past the batch entry and columns past n are clipped by the tensor map.  Versus the register path (8 LDS +
// 8 STG / RED per lane and chunk plus their address arithmetic) the exposed last epilogue of a launch is ~3x shorter.
__device__ __forceinline__ void tma_out_4d(const void* tmap, const void* src, bool reduce_add, int c0, int c1, int c2, int c3) {
  if (reduce_add)
    asm volatile("cp.reduce.async.bulk.tensor.4d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
  else
    asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void tma_out_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

// Per-column vectors of the warp's (up to four) chunks, one column per lane, fetched BEFORE the warp waits for the tile's
// MMAs (their L2 latency, ~1 k clocks under load, would otherwise sit on every chunk of the exposed epilogue); broadcast
// with shuffles when the chunk is processed.  ga / gb: the gate rows of the block's first / last row (rows_per_gate >= 32:
// a 32-row block touches at most two gate rows).
struct LeanPre {
  float ga[4], gb[4], bs[4];
};
__device__ __forceinline__ void epi_t_lean_prefetch(const int BN, const EpiParams& P, int lane, const EpiRows& R, LeanPre& A) {
  if (R.nrows <= 0) return;
  const int cend = min(BN, P.n - R.n0);
  const unsigned rpg = (unsigned)P.rows_per_gate;
  const unsigned g0 = (unsigned)R.mg0 / rpg;
  const bool two = ((unsigned)R.mg0 - g0 * rpg) + (unsigned)(R.nrows - 1) >= rpg;
  const bool fast = P.gate && rpg >= 32u;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = R.half * 32 + 64 * i;
    const int n = R.n0 + c + lane;
    const bool ok = c < cend && n < P.n;
    A.ga[i] = (fast && ok) ? __ldg(P.gate + (long long)g0 * P.gate_ld + n) : 1.f;
    A.gb[i] = (fast && ok && two) ? __ldg(P.gate + (long long)(g0 + 1u) * P.gate_ld + n) : A.ga[i];
    A.bs[i] = (P.bias && ok && R.split == 0) ? __ldg(P.bias + n) : 0.f;
  }
}

// `sc_stride` (floats): 0 = every chunk is staged in the same scratch (the bulk operation of chunk i must have read it
// before chunk i + 1 is written); > 0 = chunk i is staged at sc + i * sc_stride (the CTA's LAST tile: the operand ring is
// idle by then and serves as staging, so no chunk waits for the one before -- this is the exposed epilogue of a launch).
template <class Loader>
// gsm (nullable): the tile's gate row in shared memory, one float per tile column (gemm2.cu stages it when all of the
// CTA's 128 rows share one gate row -- always, in the DiT): broadcast LDS.128 instead of 32 shuffles per chunk.
__device__ __forceinline__ void epi_t_lean_tma(const int BN, const EpiParams& P, Loader& L, float* sc, int lane,
                                               const EpiRows& R, const void* tmap_out, const LeanPre& A, int sc_stride = 0,
                                               const float* gsm = nullptr) {
  if (R.nrows <= 0) return;                                // warp-uniform
  const int cend = min(BN, P.n - R.n0);                    // columns of this tile that exist (multiple of 8)
  const unsigned rpg = (unsigned)P.rows_per_gate;
  const unsigned g0 = (unsigned)R.mg0 / rpg, rem0 = (unsigned)R.mg0 - g0 * rpg;
  const bool two = rem0 + (unsigned)(R.nrows - 1) >= rpg;  // warp-uniform: the block straddles a gate-row boundary
  const bool hi = rem0 + (unsigned)lane >= rpg;            // this lane's row takes the second gate row
  const bool slow_gate = P.gate && rpg < 32u;              // more than two gate rows per block: per-lane loads (not the DiT)
  const bool has_bias = P.bias && R.split == 0;
  const float* grow = (slow_gate && lane < R.nrows)
                          ? P.gate + (long long)(((unsigned)R.mg0 + (unsigned)lane) / rpg) * P.gate_ld : nullptr;
  uint32_t row_addr = smem_u32(sc) + (uint32_t)lane * 128u;
  bool pending = false;
  int i = 0;
#ifdef ECHO_DEBUG
  const int dbg = g_epi_dbg;
  long long* const etr = (g_epi_trace && (threadIdx.x >> 5) == 2 && lane == 0) ? g_epi_trace + (long long)blockIdx.x * 32 : nullptr;
#else
  constexpr int dbg = 0;
#endif
  EPI_STAMP(0);
#pragma unroll 1
  for (int c = R.half * 32; c < cend; c += 64, ++i) {
    const int n = R.n0 + c;
    float v[32];
    L.load32(c, v);
    EPI_STAMP(1 + 4 * i);
    const float ga = i == 0 ? A.ga[0] : (i == 1 ? A.ga[1] : (i == 2 ? A.ga[2] : A.ga[3]));
    const float gb = i == 0 ? A.gb[0] : (i == 1 ? A.gb[1] : (i == 2 ? A.gb[2] : A.gb[3]));
    const float bs = i == 0 ? A.bs[0] : (i == 1 ? A.bs[1] : (i == 2 ? A.bs[2] : A.bs[3]));
    if (has_bias && !(dbg & 2)) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] += __shfl_sync(0xffffffffu, bs, j);
    }
    if (dbg & 2) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = v[j] * ga + bs;
    } else if (gsm) {
      const uint32_t ga_addr = smem_u32(gsm + c);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float4 g;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(g.x), "=f"(g.y), "=f"(g.z), "=f"(g.w) : "r"(ga_addr + 16u * k));
        v[4 * k] *= g.x; v[4 * k + 1] *= g.y; v[4 * k + 2] *= g.z; v[4 * k + 3] *= g.w;
      }
    } else if (P.gate && !slow_gate) {
      if (two) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const float a = __shfl_sync(0xffffffffu, ga, j), b = __shfl_sync(0xffffffffu, gb, j);
          v[j] *= hi ? b : a;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] *= __shfl_sync(0xffffffffu, ga, j);
      }
    } else if (grow) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (n + 4 * k < P.n) {
          const float4 g = __ldg(reinterpret_cast<const float4*>(grow + n) + k);
          v[4 * k] *= g.x; v[4 * k + 1] *= g.y; v[4 * k + 2] *= g.z; v[4 * k + 3] *= g.w;
        }
      }
    }
    EPI_STAMP(2 + 4 * i);
    if (pending && sc_stride == 0) {                       // the previous chunk's bulk operation has read the scratch
      if (lane == 0) tma_out_wait_read();
      __syncwarp();
    }
    if (!(dbg & 4)) {
#pragma unroll
    for (int k = 0; k < 8; ++k)
      asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(row_addr + (uint32_t)((k ^ (lane & 7)) << 4)),
                   "f"(v[4 * k]), "f"(v[4 * k + 1]), "f"(v[4 * k + 2]), "f"(v[4 * k + 3])
                   : "memory");
    } else {
      float acc = 0.f;
#pragma unroll
      for (int k = 0; k < 32; ++k) acc += v[k];
      if (acc == 12345.678f) asm volatile("st.shared.f32 [%0], %1;" ::"r"(row_addr), "f"(acc) : "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the bulk copy
    __syncwarp();
    EPI_STAMP(3 + 4 * i);
    if (lane == 0 && !(dbg & 1)) tma_out_4d(tmap_out, sc, P.red_resid != 0, n, R.mb0, R.batch_idx, R.split);
    EPI_STAMP(4 + 4 * i);
    pending = true;
    sc += sc_stride;
    row_addr += (uint32_t)sc_stride * 4u;
  }
  if (pending) {                                           // the staging memory (and, at exit, the CTA's shared memory) is free again
    if (lane == 0) tma_out_wait_read();
    __syncwarp();
  }
  EPI_STAMP(17);
}

// ---- GENERIC, row per thread, all traffic through TMA (the S1-DAC convolutions) ----------------------------------------
// v = acc + bias [+ resid];  out_f32 <- v;  bf16 out <- snake(v) (out2_bf16) or v (out_bf16).  The transposing GENERIC
// epilogue above spends ~5 k clocks per 32-column chunk on the k = 1 residual-unit convolutions (per-lane address
// arithmetic with spilled bases, two shared-memory passes, 24 predicated global accesses per lane), which makes those
// launches epilogue-bound at 60 % of the HBM rate.  Here the chunk stays in the TMEM layout (lane = row): the residual
// block arrives as a TMA load into a ping-pong buffer two chunks ahead, bias / Snake vectors are broadcast reads of a
// per-tile shared-memory row, and both outputs leave as bulk tensor stores (clipped by the tensor maps at the ends).
constexpr int EPI_GENERIC_ROW = 4;
struct RowStage {
  float* out;            // 4 KB  fp32 chunk, SWIZZLE_128B rows (this warp's slot of the transpose scratch)
  float* res[2];         // 2 x 4 KB residual chunks (TMA destinations)
  uint64_t* res_bar;     // [2]
  uint32_t* outb;        // 2 KB  bf16 chunk, linear 64-byte rows
  const float* rows;     // [3][ACC_COLS]: bias | alpha | 1 / (alpha + 1e-9) of the tile's columns
  uint32_t res_phase;    // bit b: parity res_bar[b] completes next
  int ld_rows;           // floats between the three rows
};
__device__ __forceinline__ void tma_load_res_4d(void* dst, const void* tmap, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(0)
      : "memory");
}
__device__ __forceinline__ void tma_wait_read_1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
template <class Loader>
__device__ __forceinline__ void epi_t_row(const int BN, const EpiParams& P, Loader& L, int lane, const EpiRows& R, RowStage& S,
                                          const void* tmap_out, const void* tmap_res, const void* tmap_ob) {
  if (R.nrows <= 0) return;                                // warp-uniform
  const int cend = min(BN, P.n - R.n0);
  const bool has_res = P.resid != nullptr, has_f32 = P.out_f32 != nullptr, snake = P.out2_bf16 != nullptr;
  const bool has_b = snake || P.out_bf16 != nullptr;
  const bool both = has_f32 && has_b;
  const int c_first = R.half * 32;
  if (has_res && lane == 0) {
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int cc = c_first + 64 * k;
      if (cc < cend) {
        ptx::mbar_arrive_expect_tx(&S.res_bar[k], 32 * 32 * 4);
        tma_load_res_4d(S.res[k], tmap_res, &S.res_bar[k], R.n0 + cc, R.mb0, R.batch_idx);
      }
    }
  }
  const uint32_t out_row = smem_u32(S.out) + (uint32_t)lane * 128u;
  const uint32_t outb_row = smem_u32(S.outb) + (uint32_t)lane * 64u;
  bool pend_f = false, pend_b = false;
  int i = 0;
#pragma unroll 1
  for (int c = c_first; c < cend; c += 64, ++i) {
    const int b = i & 1;
    float v[32];
    L.load32(c, v);
    const uint32_t rows_addr = smem_u32(S.rows + c);
    // v = acc + (bias + resid): the association of the transposing epilogue, so both forms give the same bits
    if (has_res) {
      ptx::mbar_wait(&S.res_bar[b], (S.res_phase >> b) & 1u);
      S.res_phase ^= 1u << b;
      const uint32_t ra = smem_u32(S.res[b]) + (uint32_t)lane * 128u;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float4 g, bs = make_float4(0.f, 0.f, 0.f, 0.f);
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(g.x), "=f"(g.y), "=f"(g.z), "=f"(g.w)
                     : "r"(ra + (uint32_t)((k ^ (lane & 7)) << 4)) : "memory");
        if (P.bias)
          asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(bs.x), "=f"(bs.y), "=f"(bs.z), "=f"(bs.w) : "r"(rows_addr + 16u * k));
        v[4 * k] += bs.x + g.x; v[4 * k + 1] += bs.y + g.y; v[4 * k + 2] += bs.z + g.z; v[4 * k + 3] += bs.w + g.w;
      }
      __syncwarp();                                        // every lane has read the buffer: refill it two chunks ahead
      if (lane == 0 && c + 128 < cend) {
        ptx::mbar_arrive_expect_tx(&S.res_bar[b], 32 * 32 * 4);
        tma_load_res_4d(S.res[b], tmap_res, &S.res_bar[b], R.n0 + c + 128, R.mb0, R.batch_idx);
      }
    }
    else if (P.bias) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float4 g;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(g.x), "=f"(g.y), "=f"(g.z), "=f"(g.w) : "r"(rows_addr + 16u * k));
        v[4 * k] += g.x; v[4 * k + 1] += g.y; v[4 * k + 2] += g.z; v[4 * k + 3] += g.w;
      }
    }
    if (has_f32) {
      if (pend_f) {                                        // the previous fp32 chunk has left the staging slot
        if (lane == 0) { if (both) tma_wait_read_1(); else tma_out_wait_read(); }
        __syncwarp();
      }
#pragma unroll
      for (int k = 0; k < 8; ++k)
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(out_row + (uint32_t)((k ^ (lane & 7)) << 4)),
                     "f"(v[4 * k]), "f"(v[4 * k + 1]), "f"(v[4 * k + 2]), "f"(v[4 * k + 3]) : "memory");
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) tma_out_4d(tmap_out, S.out, false, R.n0 + c, R.mb0, R.batch_idx, 0);
      pend_f = true;
    }
    if (has_b) {
      if (snake) {                                         // x + sin^2(alpha x) / alpha (autoencoder.py:31-41)
        const uint32_t al = rows_addr + (uint32_t)S.ld_rows * 4u, ia = rows_addr + (uint32_t)S.ld_rows * 8u;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          float4 a, r;
          asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "r"(al + 16u * k));
          asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(ia + 16u * k));
          float s0 = __sinf(a.x * v[4 * k]), s1 = __sinf(a.y * v[4 * k + 1]), s2 = __sinf(a.z * v[4 * k + 2]), s3 = __sinf(a.w * v[4 * k + 3]);
          v[4 * k] = fmaf(s0 * s0, r.x, v[4 * k]); v[4 * k + 1] = fmaf(s1 * s1, r.y, v[4 * k + 1]);
          v[4 * k + 2] = fmaf(s2 * s2, r.z, v[4 * k + 2]); v[4 * k + 3] = fmaf(s3 * s3, r.w, v[4 * k + 3]);
        }
      }
      if (pend_b) {
        if (lane == 0) { if (both) tma_wait_read_1(); else tma_out_wait_read(); }
        __syncwarp();
      }
#pragma unroll
      for (int k = 0; k < 4; ++k)
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(outb_row + 16u * k),
                     "r"(pack_bf16x2(v[8 * k], v[8 * k + 1])), "r"(pack_bf16x2(v[8 * k + 2], v[8 * k + 3])),
                     "r"(pack_bf16x2(v[8 * k + 4], v[8 * k + 5])), "r"(pack_bf16x2(v[8 * k + 6], v[8 * k + 7])) : "memory");
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) tma_out_4d(tmap_ob, S.outb, false, R.n0 + c, R.mb0, R.batch_idx, 0);
      pend_b = true;
    }
  }
  if (pend_f || pend_b) {
    if (lane == 0) tma_out_wait_read();
    __syncwarp();
  }
}

// What an epilogue can fetch before its accumulator exists (kept in registers across the wait for the MMAs).
template <int EPI> struct EpiPre {};
template <> struct EpiPre<ECHO_EPI_GENERIC> { GenAux a; };
template <> struct EpiPre<EPI_GENERIC_LEAN> { LeanPre a; };
// (SECTIONS: holding the RoPE rows of the first chunk across the wait costs 36 registers and spills; measured slower)

__device__ __forceinline__ EpiRows epi_rows(const EpiParams& P, int half, int batch_idx, int mb0, int n0, int split,
                                            long long out_off) {
  EpiRows R;
  R.batch_idx = batch_idx; R.mb0 = mb0; R.n0 = n0; R.half = half; R.split = split; R.out_off = out_off;
  R.nrows = min(32, max(0, P.rows - mb0));
  R.mg0 = (long long)batch_idx * P.rows + mb0;
  return R;
}
template <int EPI>
__device__ __forceinline__ void epilogue_prefetch_t(const int BN, const EpiParams& P, int lane, int half, int batch_idx,
                                                    int mb0, int n0, EpiPre<EPI>& pre) {
  const EpiRows R = epi_rows(P, half, batch_idx, mb0, n0, 0, 0);
  if constexpr (EPI == ECHO_EPI_GENERIC) epi_t_generic_prefetch(BN, P, lane, R, pre.a);
  else if constexpr (EPI == EPI_GENERIC_LEAN) epi_t_lean_prefetch(BN, P, lane, R, pre.a);
}
template <int EPI, class Loader>
__device__ __forceinline__ void run_epilogue_t(const int BN, const EpiParams& P, Loader& L, float* sc, int lane, int half,
                                               int batch_idx, int mb0, int n0, EpiPre<EPI>& pre, const void* tmap_out,
                                               int split = 0, long long out_off = 0, int sc_stride = 0,
                                               const float* gsm = nullptr) {
  const EpiRows R = epi_rows(P, half, batch_idx, mb0, n0, split, out_off);
  if constexpr (EPI == ECHO_EPI_GENERIC) epi_t_generic<false>(BN, P, L, sc, lane, R, pre.a);
  else if constexpr (EPI == EPI_GENERIC_LEAN) epi_t_lean_tma(BN, P, L, sc, lane, R, tmap_out, pre.a, sc_stride, gsm);
  else if constexpr (EPI == ECHO_EPI_SECTIONS) epi_t_sections(BN, P, L, sc, lane, R);
  else epi_t_swiglu(BN, P, L, sc, lane, R);
}
