// Flash-style attention over up to four KV segments walked in place (self | latent prefix | text |
// speaker for the DiT; a single causal / windowed / key-padded segment for the encoders and the
// S1-DAC post_module).  64 queries x 64 keys per step, online softmax in fp32 (exp2 domain),
// bf16 mma.sync.m16n8k16 with fp32 accumulate, cp.async double buffering, per-row extents so that
// padded or dropped (CFG) modalities cost nothing, speaker_kv_scale applied as a per-segment scalar,
// sigmoid gate fused into the output store.
#include "common.cuh"
#include "../../include/echo_b200.h"

namespace {
constexpr int ABM = 64, ABN = 64;

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  const int sz = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <int HD>
__device__ __forceinline__ uint32_t toff(int r, int chunk) { return r * (HD * 2) + ((chunk ^ (r & 7)) << 4); }

template <int HD>
__global__ void __launch_bounds__(128) attention_kernel(const __grid_constant__ echo_attn_desc D) {
  constexpr int CH = HD / 8;                 // 16-byte chunks per row
  constexpr int TILE = ABN * HD * 2;         // bytes per K (or V) tile
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint8_t s_valid[2][ABN];
  __shared__ int s_ntiles[4], s_tile0[4];
  const uint32_t sK = smem_u32(smem), sV = sK + 2 * TILE;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q0 = blockIdx.x * ABM, head = blockIdx.y, row = blockIdx.z;

  // ---- which tiles of which segments this CTA must visit -----------------------------------
  if (tid < 4) {
    int nt = 0, t0 = 0;
    if (tid < D.n_seg) {
      const echo_attn_segment& sg = D.seg[tid];
      int ext = sg.extent ? min(sg.extent[row], sg.len) : sg.len;
      if (sg.causal) {
        ext = min(ext, min(q0 + ABM, D.seq) + sg.q_offset);
        if (sg.window > 0) t0 = max(0, q0 + sg.q_offset - sg.window + 1) / ABN;
      }
      nt = max(0, (ext + ABN - 1) / ABN - t0);
    }
    s_ntiles[tid] = nt;
    s_tile0[tid] = t0;
  }

  // ---- Q tile -> shared (K stage 0 region) -> registers -----------------------------------
  {
    const bf16* qbase = (const bf16*)D.q + (long long)row * D.q_batch_stride + (long long)head * HD;
    for (int i = tid; i < ABM * CH; i += 128) {
      const int r = i / CH, c = i % CH;
      const bool ok = (q0 + r) < D.seq;
      cp_async16(sK + toff<HD>(r, c), ok ? (const void*)(qbase + (long long)(q0 + r) * D.q_row_stride + c * 8) : D.q, ok);
    }
    cp_async_commit();
    cp_async_wait<0>();
  }
  __syncthreads();
  uint32_t qf[HD / 16][4];
#pragma unroll
  for (int ks = 0; ks < HD / 16; ++ks) {
    const int r = warp * 16 + (lane & 15), c = ks * 2 + (lane >> 4);
    ldsm_x4(sK + toff<HD>(r, c), qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3]);
  }
  __syncthreads();
  const int total = s_ntiles[0] + s_ntiles[1] + s_ntiles[2] + s_ntiles[3];

  auto locate = [&](int idx, int& sgi, int& j0) {
    sgi = 0;
    while (sgi < 3 && idx >= s_ntiles[sgi]) { idx -= s_ntiles[sgi]; ++sgi; }
    j0 = (s_tile0[sgi] + idx) * ABN;
  };
  auto issue = [&](int idx, int stage) {
    int sgi, j0;
    locate(idx, sgi, j0);
    const echo_attn_segment& sg = D.seg[sgi];
    const int kvrow = sg.kv_batch_mod > 0 ? row % sg.kv_batch_mod : row;
    const bf16* kb = (const bf16*)sg.k + (long long)kvrow * sg.batch_stride + (long long)head * HD;
    const bf16* vb = (const bf16*)sg.v + (long long)kvrow * sg.batch_stride + (long long)head * HD;
    for (int i = tid; i < ABN * CH; i += 128) {
      const int r = i / CH, c = i % CH;
      const bool ok = (j0 + r) < sg.len;
      const long long off = (long long)(j0 + r) * sg.row_stride + c * 8;
      cp_async16(sK + stage * TILE + toff<HD>(r, c), ok ? (const void*)(kb + off) : sg.k, ok);
      cp_async16(sV + stage * TILE + toff<HD>(r, c), ok ? (const void*)(vb + off) : sg.v, ok);
    }
    if (tid < ABN) {
      const int j = j0 + tid;
      uint8_t ok = j < sg.len;
      if (ok && sg.mask) ok = sg.mask[(long long)row * sg.mask_stride + j] != 0;
      s_valid[stage][tid] = ok;
    }
    cp_async_commit();
  };

  float oacc[HD / 8][4];
#pragma unroll
  for (int i = 0; i < HD / 8; ++i) oacc[i][0] = oacc[i][1] = oacc[i][2] = oacc[i][3] = 0.f;
  float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};
  const int qi0 = q0 + warp * 16 + (lane >> 2);     // query index of accumulator rows 0 (and +8)

  if (total > 0) issue(0, 0);
  for (int it = 0; it < total; ++it) {
    const int stage = it & 1;
    if (it + 1 < total) { issue(it + 1, stage ^ 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    int sgi, j0;
    locate(it, sgi, j0);
    const echo_attn_segment& sg = D.seg[sgi];
    const float f = sg.kv_scale ? sg.kv_scale[0] : 1.f;
    const float sl2 = D.softmax_scale * 1.4426950408889634f * f;
    const uint32_t kt = sK + stage * TILE, vt = sV + stage * TILE;

    float sacc[ABN / 8][4];
#pragma unroll
    for (int i = 0; i < ABN / 8; ++i) sacc[i][0] = sacc[i][1] = sacc[i][2] = sacc[i][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < HD / 16; ++ks) {
#pragma unroll
      for (int nj = 0; nj < ABN / 16; ++nj) {
        uint32_t b0, b1, b2, b3;
        const int r = nj * 16 + (lane & 7) + ((lane >> 4) << 3), c = ks * 2 + ((lane >> 3) & 1);
        ldsm_x4(kt + toff<HD>(r, c), b0, b1, b2, b3);
        mma16816(sacc[nj * 2], qf[ks], b0, b1);
        mma16816(sacc[nj * 2 + 1], qf[ks], b2, b3);
      }
    }
    // mask + online softmax (rows qi0 and qi0 + 8)
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int ni = 0; ni < ABN / 8; ++ni) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int col = ni * 8 + (lane & 3) * 2 + (e & 1);
        const int qi = qi0 + (e >> 1) * 8;
        bool ok = s_valid[stage][col] != 0;
        if (sg.causal) {
          const int j = j0 + col, qk = qi + sg.q_offset;
          ok = ok && (j <= qk) && (sg.window <= 0 || j > qk - sg.window);
        }
        const float s = ok ? sacc[ni][e] * sl2 : -INFINITY;
        sacc[ni][e] = s;
        mx[e >> 1] = fmaxf(mx[e >> 1], s);
      }
    }
    float alpha[2], muse[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 1));
      mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 2));
      const float mnew = fmaxf(mrow[h], mx[h]);
      muse[h] = (mnew == -INFINITY) ? 0.f : mnew;
      alpha[h] = (mrow[h] == -INFINITY) ? 0.f : exp2f(mrow[h] - mnew);
      mrow[h] = mnew;
      lrow[h] *= alpha[h];
    }
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) {
      oacc[i][0] *= alpha[0]; oacc[i][1] *= alpha[0];
      oacc[i][2] *= alpha[1]; oacc[i][3] *= alpha[1];
    }
    uint32_t pf[ABN / 16][4];
#pragma unroll
    for (int ni = 0; ni < ABN / 8; ++ni) {
      const float p0 = exp2f(sacc[ni][0] - muse[0]), p1 = exp2f(sacc[ni][1] - muse[0]);
      const float p2 = exp2f(sacc[ni][2] - muse[1]), p3 = exp2f(sacc[ni][3] - muse[1]);
      lrow[0] += p0 + p1;
      lrow[1] += p2 + p3;
      pf[ni >> 1][(ni & 1) * 2 + 0] = pack_bf16x2(p0 * f, p1 * f);
      pf[ni >> 1][(ni & 1) * 2 + 1] = pack_bf16x2(p2 * f, p3 * f);
    }
#pragma unroll
    for (int kk = 0; kk < ABN / 16; ++kk) {
#pragma unroll
      for (int nd = 0; nd < HD / 16; ++nd) {
        uint32_t b0, b1, b2, b3;
        const int r = kk * 16 + (lane & 7) + (((lane >> 3) & 1) << 3), c = nd * 2 + (lane >> 4);
        ldsm_x4_t(vt + toff<HD>(r, c), b0, b1, b2, b3);
        mma16816(oacc[nd * 2], pf[kk], b0, b1);
        mma16816(oacc[nd * 2 + 1], pf[kk], b2, b3);
      }
    }
    __syncthreads();
  }

  // ---- normalise, gate, store -----------------------------------------------------------------
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    lrow[h] += __shfl_xor_sync(0xffffffffu, lrow[h], 1);
    lrow[h] += __shfl_xor_sync(0xffffffffu, lrow[h], 2);
  }
  const int width = D.heads * HD;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int qi = qi0 + h * 8;
    if (qi >= D.seq) continue;
    const float inv = lrow[h] > 0.f ? 1.f / lrow[h] : 0.f;
    const long long o = ((long long)row * D.seq + qi) * width + (long long)head * HD;
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) {
      const int col = i * 8 + (lane & 3) * 2;
      float a = oacc[i][h * 2] * inv, b = oacc[i][h * 2 + 1] * inv;
      if (D.gate) {
        const float2 g = unpack_bf16x2(*reinterpret_cast<const uint32_t*>((const bf16*)D.gate + o + col));
        a *= g.x; b *= g.y;
      }
      *reinterpret_cast<uint32_t*>((bf16*)D.out + o + col) = pack_bf16x2(a, b);
    }
  }
}
}  // namespace

int echo_attention_tc(const echo_attn_desc* d, cudaStream_t st);   // attention2.cu (tcgen05 kernel)

extern "C" int echo_attention(const echo_attn_desc* d, echo_stream_t stream) {
  ECHO_CHECK_ARG(d && d->q && d->out, "attention: null argument");
  ECHO_CHECK_ARG(d->head_dim == 64 || d->head_dim == 128, "attention: head_dim must be 64 or 128");
  ECHO_CHECK_ARG(d->n_seg >= 1 && d->n_seg <= 4, "attention: 1..4 segments");
  ECHO_CHECK_ARG(d->rows > 0 && d->seq > 0 && d->heads > 0, "attention: bad shape");
  ECHO_CHECK_ARG(d->q_row_stride % 8 == 0 && d->q_batch_stride % 8 == 0, "attention: q strides must be multiples of 8");
  for (int i = 0; i < d->n_seg; ++i) {
    ECHO_CHECK_ARG(d->seg[i].len == 0 || (d->seg[i].k && d->seg[i].v), "attention: segment %d has no K/V", i);
    ECHO_CHECK_ARG(d->seg[i].row_stride % 8 == 0 && d->seg[i].batch_stride % 8 == 0, "attention: segment strides %%8");
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (d->impl == 0) {   // tcgen05 kernel when the shape is in its domain (head_dim 128, no sliding window)
    const int rc = echo_attention_tc(d, st);
    if (rc != ECHO_ERR_UNSUPPORTED) return rc;
  }
  dim3 grid(ceil_div(d->seq, ABM), d->heads, d->rows);
  if (d->head_dim == 128) {
    const int smem = 4 * ABN * 128 * 2;
    static bool set = false;
    if (!set) { ECHO_CUDA(cudaFuncSetAttribute(attention_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); set = true; }
    attention_kernel<128><<<grid, 128, smem, st>>>(*d);
  } else {
    const int smem = 4 * ABN * 64 * 2;
    attention_kernel<64><<<grid, 128, smem, st>>>(*d);
  }
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}
