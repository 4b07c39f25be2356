// fp32-class arithmetic for the S1-DAC ENCODE path (DAC.encode_zq, autoencoder.py:1080-1126).
//
// The encode path ends in a nearest-code search (VectorQuantize.decode_latents, autoencoder.py:145-157): an integer
// decision, which plain bf16 tensor-core operands flip on 1-8 % of the frames.  Here every contraction of the path runs
// on the same tcgen05 GEMM kernel with BOTH operands split into two bf16 terms (x = hi + lo, 16 mantissa bits) and the
// three significant products accumulated in fp32:  x w ~= hi w_hi + lo w_hi + hi w_lo  (the dropped lo w_lo term is
// 2^-16 relative).  The split operand is laid out along K as [hi | lo | hi] against weights packed [w_hi | w_hi | w_lo]
// per tap, so the GEMM kernel needs no new mode -- K is three times longer.  Activations between GEMMs stay fp32 in
// HBM; this file holds the elementwise pieces: the split (with Snake / SwiGLU fused in front), fp32 causal-window
// attention with rotary embedding for the two small transformers, and the ConvNeXt depthwise conv + LayerNorm in fp32.
#include <math.h>
#include "common.cuh"
#include "../../include/echo_b200.h"

namespace {

// ---- split: fp32 [rows, C] -> bf16 [rows, 3C] = [hi | lo | hi] ------------------------------------------------------
// mode 0: y = x;  1: y = snake(x, alpha) (autoencoder.py:96-109, precise sinf);  2: y = silu(x[:, c]) * x[:, C + c]
// (SwiGLU of an un-interleaved [w1 x | w3 x] row, autoencoder.py:717)
__global__ void __launch_bounds__(256) split_bf16x3_kernel(const float* __restrict__ x, long long ldx, long long rows, int c_len,
                                                           int mode, const float* __restrict__ alpha, bf16* __restrict__ out,
                                                           long long ldo) {
  const long long n4 = rows * (long long)(c_len >> 2);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / (c_len >> 2);
    const int c = (int)(i - r * (c_len >> 2)) << 2;
    float4 v = *reinterpret_cast<const float4*>(x + r * ldx + c);
    if (mode == 1) {
      const float4 a = *reinterpret_cast<const float4*>(alpha + c);
      float s;
      s = sinf(a.x * v.x); v.x = v.x + s * s / (a.x + 1e-9f);
      s = sinf(a.y * v.y); v.y = v.y + s * s / (a.y + 1e-9f);
      s = sinf(a.z * v.z); v.z = v.z + s * s / (a.z + 1e-9f);
      s = sinf(a.w * v.w); v.w = v.w + s * s / (a.w + 1e-9f);
    } else if (mode == 2) {
      const float4 g = *reinterpret_cast<const float4*>(x + r * ldx + c_len + c);
      v.x = v.x / (1.f + expf(-v.x)) * g.x; v.y = v.y / (1.f + expf(-v.y)) * g.y;
      v.z = v.z / (1.f + expf(-v.z)) * g.z; v.w = v.w / (1.f + expf(-v.w)) * g.w;
    }
    const bf16 h0 = __float2bfloat16(v.x), h1 = __float2bfloat16(v.y), h2 = __float2bfloat16(v.z), h3 = __float2bfloat16(v.w);
    const float l0 = v.x - __bfloat162float(h0), l1 = v.y - __bfloat162float(h1), l2 = v.z - __bfloat162float(h2),
                l3 = v.w - __bfloat162float(h3);
    uint2 hi, lo;
    hi.x = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
    hi.y = (uint32_t)__bfloat16_as_ushort(h2) | ((uint32_t)__bfloat16_as_ushort(h3) << 16);
    lo.x = pack_bf16x2(l0, l1);
    lo.y = pack_bf16x2(l2, l3);
    bf16* o = out + r * ldo + c;
    *reinterpret_cast<uint2*>(o) = hi;
    *reinterpret_cast<uint2*>(o + c_len) = lo;
    *reinterpret_cast<uint2*>(o + 2 * c_len) = hi;
  }
}

// ---- rotary embedding in place on fp32 rows (AE apply_rotary_emb, autoencoder.py:816-826: interleaved pairs, every head;
//      the table is the reference's bf16-rounded freqs_cis buffer, :805-813) ----------------------------------------
__global__ void rope_f32_kernel(float* __restrict__ x, long long ldx, long long rows, int heads, int head_dim,
                                const float* __restrict__ cos_t, const float* __restrict__ sin_t, int rows_per_seq,
                                int pos_offset) {
  const int half = head_dim >> 1;
  const long long n = rows * heads * half;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int p = (int)(i % half);
    const long long rh = i / half;
    const int h = (int)(rh % heads);
    const long long r = rh / heads;
    const int pos = pos_offset + (int)(r % rows_per_seq);
    const float co = cos_t[(long long)pos * half + p], si = sin_t[(long long)pos * half + p];
    float2* e = reinterpret_cast<float2*>(x + r * ldx + (long long)h * head_dim + 2 * p);
    const float2 v = *e;
    *e = make_float2(v.x * co - v.y * si, v.y * co + v.x * si);
  }
}

// ---- fp32 causal sliding-window attention (WindowLimitedTransformer, autoencoder.py:762-773: key j visible to query i
//      iff i - window < j <= i), head_dim 64.  q / k / v: fp32 rows [B*T, ld] with head h at columns h*64 (q, k already
//      rotated); out fp32 [B*T, H*64].  One CTA = 16 consecutive queries of one (batch, head); 4 warps x 4 queries; keys
//      in shared-memory tiles of 64; lane l scores keys l and l+32 of the tile and owns output dims l and l+32.
constexpr int AQ = 16, AKT = 64, AHD = 64;
__global__ void __launch_bounds__(128) attention_f32_kernel(const float* __restrict__ q, const float* __restrict__ k,
                                                            const float* __restrict__ v, long long ld, int t_len, int window,
                                                            float scale, float* __restrict__ out, long long ldo) {
  __shared__ float sk[AKT][AHD + 1];
  __shared__ float sv[AKT][AHD];
  __shared__ float sq[AQ][AHD];
  const int q0 = blockIdx.x * AQ, h = blockIdx.y, b = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row0 = (long long)b * t_len;
  const int col = h * AHD;
  for (int i = threadIdx.x; i < AQ * AHD; i += 128) {
    const int r = i / AHD, c = i % AHD;
    sq[r][c] = (q0 + r < t_len) ? q[(row0 + q0 + r) * ld + col + c] * scale : 0.f;
  }
  float m[4], l[4], o0[4], o1[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) { m[j] = -INFINITY; l[j] = 0.f; o0[j] = 0.f; o1[j] = 0.f; }
  const int q_last = min(q0 + AQ, t_len) - 1;
  const int k_begin = max(0, q0 - window + 1), k_end = q_last + 1;       // keys any query of the CTA can see
  for (int kt = (k_begin / AKT) * AKT; kt < k_end; kt += AKT) {
    __syncthreads();
    for (int i = threadIdx.x; i < AKT * AHD; i += 128) {
      const int r = i / AHD, c = i % AHD;
      const bool ok = kt + r < t_len;
      sk[r][c] = ok ? k[(row0 + kt + r) * ld + col + c] : 0.f;
      sv[r][c] = ok ? v[(row0 + kt + r) * ld + col + c] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int qi = q0 + warp * 4 + j;
      if (qi >= t_len) continue;                                          // warp-uniform
      const int lo = qi - window + 1;
      float s0 = 0.f, s1 = 0.f;
      const float* qr = sq[warp * 4 + j];
#pragma unroll 16
      for (int c = 0; c < AHD; ++c) {
        const float qv = qr[c];
        s0 = fmaf(qv, sk[lane][c], s0);
        s1 = fmaf(qv, sk[lane + 32][c], s1);
      }
      const int k0 = kt + lane, k1 = kt + lane + 32;
      const bool v0 = k0 <= qi && k0 >= lo, v1 = k1 <= qi && k1 >= lo;
      s0 = v0 ? s0 : -INFINITY;
      s1 = v1 ? s1 : -INFINITY;
      const float tmax = warp_max(fmaxf(s0, s1));
      if (tmax == -INFINITY) continue;                                    // no visible key in this tile (warp-uniform)
      const float mn = fmaxf(m[j], tmax);
      const float corr = expf(m[j] - mn);                                 // exp(-inf) = 0 on the first tile
      const float p0 = v0 ? expf(s0 - mn) : 0.f, p1 = v1 ? expf(s1 - mn) : 0.f;
      l[j] = l[j] * corr + warp_sum(p0 + p1);
      float a0 = o0[j] * corr, a1 = o1[j] * corr;
#pragma unroll 8
      for (int kk = 0; kk < 32; ++kk) {
        const float pa = __shfl_sync(0xffffffffu, p0, kk), pb = __shfl_sync(0xffffffffu, p1, kk);
        a0 = fmaf(pa, sv[kk][lane], a0);       a1 = fmaf(pa, sv[kk][lane + 32], a1);
        a0 = fmaf(pb, sv[kk + 32][lane], a0);  a1 = fmaf(pb, sv[kk + 32][lane + 32], a1);
      }
      o0[j] = a0; o1[j] = a1; m[j] = mn;
    }
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int qi = q0 + warp * 4 + j;
    if (qi >= t_len) continue;
    const float inv = 1.f / l[j];
    float* orow = out + (row0 + qi) * ldo + col;
    orow[lane] = o0[j] * inv;
    orow[lane + 32] = o1[j] * inv;
  }
}

// ---- ConvNeXt front in fp32 (autoencoder.py:362-364): causal depthwise conv k = 7, then LayerNorm over channels ----
__global__ void dwconv_ln_f32_kernel(const float* __restrict__ x, int t_len, int c_len, const float* __restrict__ w,
                                     const float* __restrict__ b, const float* __restrict__ ln_w,
                                     const float* __restrict__ ln_b, float eps, float* __restrict__ y) {
  const int t = blockIdx.x, bt = blockIdx.y;
  const float* xb = x + (long long)bt * t_len * c_len;
  __shared__ float red[32];
  __shared__ float stat;
  const int nper = (c_len + blockDim.x - 1) / blockDim.x;
  float vals[4];
  float sum = 0.f;
  for (int i = 0; i < nper; ++i) {
    const int c = threadIdx.x + i * blockDim.x;
    float acc = 0.f;
    if (c < c_len) {
      acc = b[c];
#pragma unroll
      for (int kk = 0; kk < 7; ++kk) {
        const int tt = t - 6 + kk;
        if (tt >= 0) acc = fmaf(w[c * 7 + kk], xb[(long long)tt * c_len + c], acc);
      }
      sum += acc;
    }
    vals[i] = acc;
  }
  auto block_sum = [&](float val) {
    val = warp_sum(val);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = val;
    __syncthreads();
    if (threadIdx.x == 0) {
      float s = 0.f;
      for (int i = 0; i < (int)(blockDim.x >> 5); ++i) s += red[i];
      stat = s;
    }
    __syncthreads();
    const float r = stat;
    __syncthreads();
    return r;
  };
  const float mean = block_sum(sum) / (float)c_len;
  float sqs = 0.f;
  for (int i = 0; i < nper; ++i) {
    const int c = threadIdx.x + i * blockDim.x;
    if (c < c_len) { const float dlt = vals[i] - mean; sqs += dlt * dlt; }
  }
  const float rstd = 1.f / sqrtf(block_sum(sqs) / (float)c_len + eps);
  float* yr = y + ((long long)bt * t_len + t) * c_len;
  for (int i = 0; i < nper; ++i) {
    const int c = threadIdx.x + i * blockDim.x;
    if (c < c_len) yr[c] = (vals[i] - mean) * rstd * ln_w[c] + ln_b[c];
  }
}
}  // namespace

extern "C" int echo_split_bf16x3(const float* x, int64_t ldx, int64_t rows, int c, int mode, const float* alpha, void* out,
                                 int64_t ldo, echo_stream_t stream) {
  ECHO_CHECK_ARG(x && out && rows > 0 && c > 0 && c % 4 == 0 && ldx % 4 == 0 && ldo % 4 == 0, "split: bad shape (c, ldx, ldo %% 4)");
  ECHO_CHECK_ARG(mode >= 0 && mode <= 2 && (mode != 1 || alpha), "split: mode 0 plain / 1 snake (alpha) / 2 swiglu");
  ECHO_CHECK_ARG(((uintptr_t)x % 16) == 0 && ((uintptr_t)out % 8) == 0 && ((uintptr_t)alpha % 16) == 0, "split: misaligned pointer");
  const long long n4 = rows * (long long)(c / 4);
  const int grid = (int)((n4 + 255) / 256 < 148 * 16 ? (n4 + 255) / 256 : 148 * 16);
  split_bf16x3_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, ldx, rows, c, mode, alpha, (bf16*)out, ldo);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

extern "C" int echo_rope_f32(float* x, int64_t ldx, int64_t rows, int heads, int head_dim, const float* cos_t,
                             const float* sin_t, int rope_len, int rows_per_seq, int pos_offset, echo_stream_t stream) {
  ECHO_CHECK_ARG(x && cos_t && sin_t && rows > 0 && heads > 0 && head_dim > 0 && head_dim % 2 == 0 && ldx % 2 == 0,
                 "rope_f32: bad argument");
  ECHO_CHECK_ARG(rows_per_seq > 0 && pos_offset >= 0 && pos_offset + (rows_per_seq < rows ? rows_per_seq : (int)rows) <= rope_len,
                 "rope_f32: rope table too short");
  const long long n = rows * heads * (head_dim / 2);
  const int grid = (int)((n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16);
  rope_f32_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, ldx, rows, heads, head_dim, cos_t, sin_t, rows_per_seq, pos_offset);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

extern "C" int echo_attention_window_f32(const float* q, const float* k, const float* v, int64_t ld, int batch, int t,
                                         int heads, int head_dim, int window, float softmax_scale, float* out, int64_t ldo,
                                         echo_stream_t stream) {
  ECHO_CHECK_ARG(q && k && v && out && batch > 0 && t > 0 && heads > 0 && window > 0, "attention_f32: bad argument");
  ECHO_CHECK_ARG(head_dim == AHD, "attention_f32: head_dim must be 64 (the S1-DAC transformers)");
  attention_f32_kernel<<<dim3((t + AQ - 1) / AQ, heads, batch), 128, 0, (cudaStream_t)stream>>>(q, k, v, ld, t, window,
                                                                                                softmax_scale, out, ldo);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

extern "C" int echo_dwconv_layernorm_f32(const float* x, int batch, int t, int c, const float* w_dw, const float* b_dw,
                                         const float* ln_w, const float* ln_b, float eps, float* y, echo_stream_t stream) {
  ECHO_CHECK_ARG(x && w_dw && b_dw && ln_w && ln_b && y && batch > 0 && t > 0, "dwconv_ln_f32: bad argument");
  ECHO_CHECK_ARG(c > 0 && c <= 1024, "dwconv_ln_f32: c must be <= 1024");
  const int threads = c >= 256 ? 256 : ((c + 31) / 32) * 32;
  dwconv_ln_f32_kernel<<<dim3(t, batch), threads, 0, (cudaStream_t)stream>>>(x, t, c, w_dw, b_dw, ln_w, ln_b, eps, y);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}
