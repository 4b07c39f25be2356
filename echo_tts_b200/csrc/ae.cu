// S1-DAC decode helpers that are not GEMM-shaped (HBM-bound), time-major [B, T, C] activations,
// plus the flattening-point search.
#include "common.cuh"
#include "../../include/echo_b200.h"

namespace {

// ConvNeXt front (autoencoder.py:362-364): causal depthwise conv k=7 then LayerNorm over channels.
// One CTA per time step, one thread per channel (coalesced across channels), two-pass variance.
// left_ctx > 0 (streaming decode, batch 1): that many rows of the previous call's input sit in memory before row 0.
__global__ void dwconv_ln_kernel(const float* __restrict__ x, int t_len, int c_len, const float* __restrict__ w,
                                 const float* __restrict__ b, const float* __restrict__ ln_w,
                                 const float* __restrict__ ln_b, float eps, bf16* __restrict__ y, int left_ctx) {
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
      for (int k = 0; k < 7; ++k) {
        const int tt = t - 6 + k;
        if (tt >= -left_ctx) acc = fmaf(w[c * 7 + k], xb[(long long)tt * c_len + c], acc);
      }
      sum += acc;
    }
    vals[i] = acc;
  }
  auto block_sum = [&](float v) {
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
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
  float sq = 0.f;
  for (int i = 0; i < nper; ++i) {
    const int c = threadIdx.x + i * blockDim.x;
    if (c < c_len) { const float dlt = vals[i] - mean; sq += dlt * dlt; }
  }
  const float rstd = rsqrtf(block_sum(sq) / (float)c_len + eps);
  bf16* yr = y + ((long long)bt * t_len + t) * c_len;
  for (int i = 0; i < nper; ++i) {
    const int c = threadIdx.x + i * blockDim.x;
    if (c < c_len) yr[c] = __float2bfloat16((vals[i] - mean) * rstd * ln_w[c] + ln_b[c]);
  }
}

// Last stage (autoencoder.py:994): causal conv k=7, C -> 1, + tanh on Snake'd bf16 input.  HBM-bound (C*2 bytes in, 4 out
// per sample).  One warp per run of 32*RUNS outputs, channel PAIRS across lanes (lane l owns pairs l, l+32, ...: 4-byte
// loads, a warp load covers 128 contiguous bytes of a row): every input row is read ONCE from global (no shared memory)
// and scattered into the <= 7 outputs it feeds; the 38 rows of a run are requested in batches of 10 (19) that are each
// wholly in flight before the first use (1.9 KB per warp: the kernel is a pure stream, and bytes in flight are what it
// needs); 32 per-lane partial outputs live in registers and one 31-shuffle transpose-reduction per 32 outputs leaves
// output t0 + l in lane l for a coalesced store.  left_ctx as in dwconv_ln_kernel.  c_len must be even.
template <int NP>
__global__ void __launch_bounds__(256, 2) conv_out_tanh_kernel(const bf16* __restrict__ x, int t_len, int c_len,
                                                            const float* __restrict__ w, float bias,
                                                            float* __restrict__ wav, int left_ctx, int runs) {
  const int lane = threadIdx.x & 31, bt = blockIdx.y;
  const long long warp_id = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const bf16* xb = x + (long long)bt * t_len * c_len;
  const int npairs = c_len >> 1;
  float wk[7][NP][2];
#pragma unroll
  for (int k = 0; k < 7; ++k)
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      const int pr = lane + 32 * j;
      wk[k][j][0] = pr < npairs ? __ldg(w + k * c_len + 2 * pr) : 0.f;
      wk[k][j][1] = pr < npairs ? __ldg(w + k * c_len + 2 * pr + 1) : 0.f;
    }
  for (int run = 0; run < runs; ++run) {
    const long long t0 = (warp_id * runs + run) * 32;
    if (t0 >= t_len) return;
    float acc[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) acc[i] = 0.f;
    constexpr int BR = NP >= 2 ? 10 : 19;      // rows per batch (registers): 4 or 2 batches cover the run's 38 rows
#pragma unroll
    for (int r0 = 0; r0 < 38; r0 += BR) {
      uint32_t xr[BR][NP];
#pragma unroll
      for (int i = 0; i < BR; ++i) {
        const long long r = t0 - 6 + r0 + i;
#pragma unroll
        for (int j = 0; j < NP; ++j) {
          const int pr = lane + 32 * j;
          xr[i][j] = (r0 + i < 38 && r >= -left_ctx && r < t_len && pr < npairs)
                         ? __ldg(reinterpret_cast<const uint32_t*>(xb + r * c_len) + pr) : 0u;
        }
      }
#pragma unroll
      for (int i = 0; i < BR; ++i) {
        const int rl = r0 + i;
#pragma unroll
        for (int k = 0; k < 7; ++k) {
          const int tl = rl - k;               // row r = t - 6 + k of output t = t0 + tl
          if (tl >= 0 && tl < 32) {
#pragma unroll
            for (int j = 0; j < NP; ++j) {
              const float2 v = unpack_bf16x2(xr[i][j]);
              acc[tl] = fmaf(wk[k][j][1], v.y, fmaf(wk[k][j][0], v.x, acc[tl]));
            }
          }
        }
      }
    }
    // transpose-reduce: afterwards acc[0] of lane l = sum over lanes of acc[l]
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
      const bool up = (lane & s) != 0;
#pragma unroll
      for (int i = 0; i < s; ++i) {
        const float send = up ? acc[i] : acc[i + s];
        const float keep = up ? acc[i + s] : acc[i];
        acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
      }
    }
    if (t0 + lane < t_len) wav[(long long)bt * t_len + t0 + lane] = tanhf(acc[0] + bias);
  }
}

__global__ void cast_snake_kernel(const float* __restrict__ x, long long n, int c_len, const float* __restrict__ alpha,
                                  bf16* __restrict__ y) {
  const long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= n) return;
  const float4 v = *reinterpret_cast<const float4*>(x + i);
  float o[4] = {v.x, v.y, v.z, v.w};
  if (alpha) {
    const int c = (int)(i % c_len);
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = snakef_(o[j], alpha[c + j]);
  }
  *reinterpret_cast<uint2*>(y + i) = make_uint2(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]));
}

// ---- S1-DAC encode helpers ----------------------------------------------------------------------------
// Encoder's first layer (autoencoder.py:917): causal conv k=7 from the mono waveform to C channels, fused with the
// first ResidualUnit's Snake: x fp32 [B, T, C] (residual stream) and snake(x) bf16 (operand of the next conv).
// HBM-bound on the stores: one thread per (t, 4 channels), the 7 input samples come from L1.
__global__ void conv_in_snake_kernel(const float* __restrict__ audio, int t_len, int c_len, const float* __restrict__ w,
                                     const float* __restrict__ bias, const float* __restrict__ alpha,
                                     float* __restrict__ x, bf16* __restrict__ s) {
  const int cq = c_len >> 2;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;     // over t * cq
  const int bt = blockIdx.y;
  if (i >= (long long)t_len * cq) return;
  const int t = (int)(i / cq), c = (int)(i % cq) * 4;
  const float* a = audio + (long long)bt * t_len;
  float in[7];
#pragma unroll
  for (int k = 0; k < 7; ++k) in[k] = (t - 6 + k >= 0) ? __ldg(a + t - 6 + k) : 0.f;
  float o[4], sn[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float acc = bias[c + j];
#pragma unroll
    for (int k = 0; k < 7; ++k) acc = fmaf(w[(c + j) * 7 + k], in[k], acc);
    o[j] = acc;
    sn[j] = snakef_(acc, alpha[c + j]);
  }
  const long long off = ((long long)bt * t_len + t) * c_len + c;
  *reinterpret_cast<float4*>(x + off) = make_float4(o[0], o[1], o[2], o[3]);
  *reinterpret_cast<uint2*>(s + off) = make_uint2(pack_bf16x2(sn[0], sn[1]), pack_bf16x2(sn[2], sn[3]));
}

__global__ void snake_bf16_kernel(const bf16* __restrict__ x, long long n, int c_len, const float* __restrict__ alpha,
                                  bf16* __restrict__ y) {
  const long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= n) return;
  const uint2 u = *reinterpret_cast<const uint2*>(x + i);
  const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y);
  float o[4] = {a.x, a.y, b.x, b.y};
  const int c = (int)(i % c_len);
#pragma unroll
  for (int j = 0; j < 4; ++j) o[j] = snakef_(o[j], alpha[c + j]);
  *reinterpret_cast<uint2*>(y + i) = make_uint2(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]));
}

// VectorQuantize.decode_latents (autoencoder.py:145-157) in fp32: one warp per row; the row is l2-normalised, the
// codebook arrives normalised (cbn), dist = |e|^2 - 2 e.c + |c|^2 evaluated in the reference's order; smallest
// distance wins, smallest index on exact ties (torch.max of -dist).  Also gathers the RAW code vector (embed_code).
template <int DIM>
__global__ void vq_nearest_kernel(const float* __restrict__ e, int m, const float* __restrict__ cbn,
                                  const float* __restrict__ cb, int n, int* __restrict__ idx_out,
                                  float* __restrict__ code_out) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= m) return;
  float v[DIM], ss = 0.f;
#pragma unroll
  for (int j = 0; j < DIM; ++j) { v[j] = e[(long long)row * DIM + j]; ss = fmaf(v[j], v[j], ss); }
  const float inv = 1.f / fmaxf(sqrtf(ss), 1e-12f);
  float e2 = 0.f;
#pragma unroll
  for (int j = 0; j < DIM; ++j) { v[j] *= inv; e2 = fmaf(v[j], v[j], e2); }
  float best = INFINITY;
  int bi = 0x7fffffff;
  for (int k = lane; k < n; k += 32) {
    float dot = 0.f, c2 = 0.f;
#pragma unroll
    for (int j = 0; j < DIM; ++j) {
      const float c = __ldg(cbn + (long long)k * DIM + j);
      dot = fmaf(v[j], c, dot);
      c2 = fmaf(c, c, c2);
    }
    const float dist = (e2 - 2.f * dot) + c2;
    if (dist < best) { best = dist; bi = k; }           // k ascending per lane: first minimum kept
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ob < best || (ob == best && oi < bi)) { best = ob; bi = oi; }
  }
  if (lane == 0) idx_out[row] = bi;
  if (code_out && lane < DIM) code_out[(long long)row * DIM + lane] = cb[(long long)bi * DIM + lane];
}

// find_flattening_point (inference.py:315-330).  Per-row (sum, sum of squares) in fp64, then every
// window start i checks unbiased std < thr and |mean - target| < 0.1; the smallest hit wins.
__global__ void flattening_kernel(const float* __restrict__ x, int len, int width, int window, float thr, float target,
                                  int* __restrict__ out) {
  extern __shared__ double sd[];           // [len + window][2]
  const int padded = len + window;
  for (int r = threadIdx.x >> 5; r < padded; r += blockDim.x >> 5) {
    double s = 0.0, q = 0.0;
    if (r < len)
      for (int c = threadIdx.x & 31; c < width; c += 32) {
        const double v = (double)x[(long long)r * width + c];
        s += v; q += v * v;
      }
    for (int o = 16; o > 0; o >>= 1) {
      s += __shfl_xor_sync(0xffffffffu, s, o);
      q += __shfl_xor_sync(0xffffffffu, q, o);
    }
    if ((threadIdx.x & 31) == 0) { sd[2 * r] = s; sd[2 * r + 1] = q; }
  }
  __shared__ int best;
  if (threadIdx.x == 0) best = len;
  __syncthreads();
  const double n = (double)window * width;
  for (int i = threadIdx.x; i < len; i += blockDim.x) {
    double s = 0.0, q = 0.0;
    for (int k = 0; k < window; ++k) { s += sd[2 * (i + k)]; q += sd[2 * (i + k) + 1]; }
    const double mean = s / n;
    const double var = (q - s * mean) / (n - 1.0);
    const float stdv = (float)sqrt(var > 0.0 ? var : 0.0);
    if (stdv < thr && fabsf((float)mean - target) < 0.1f) atomicMin(&best, i);
  }
  __syncthreads();
  if (threadIdx.x == 0) out[0] = best;
}
// ---- audio front-end (load_audio, inference.py:141-149) ------------------------------------------------------
// Channel mean + polyphase windowed-sinc resampling (the algorithm of torchaudio.functional.resample, which the reference
// calls at inference.py:146): output j = f * up + p is sum_k table[k][p] * mono[f * down + k - width], mono = mean over
// channels, samples outside [0, n) are zero.  The table arrives transposed ([taps][up]) so that the `up` phases of one
// frame read it coalesced; the input window of a frame is shared by all its phases (L1 broadcast).  HBM-bound on the
// output stream.  Also folds the peak search of inference.py:147 (max |y|) into the same pass: non-negative floats order
// like their bit patterns, so one atomicMax per block on the uint32 view is exact.
__global__ void __launch_bounds__(256) resample_mono_kernel(const float* __restrict__ x, int channels, long long n,
                                                            long long ldx, int down, int up, int width, int taps,
                                                            const float* __restrict__ table_t,
                                                            const int* __restrict__ tap_lo, const int* __restrict__ tap_hi,
                                                            float* __restrict__ y, long long n_out,
                                                            unsigned int* __restrict__ absmax_bits) {
  const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  float acc = 0.f;
  if (j < n_out) {
    const long long f = j / up;
    const int p = (int)(j - f * up);
    const long long i0 = f * down - width;
    const float inv_c = 1.f / (float)channels;
    const int k0 = tap_lo ? __ldg(tap_lo + p) : 0, k1 = tap_hi ? __ldg(tap_hi + p) : taps;
    for (int k = k0; k < k1; ++k) {
      const long long i = i0 + k;
      if (i < 0 || i >= n) continue;
      float m = __ldg(x + i);
      for (int c = 1; c < channels; ++c) m += __ldg(x + c * ldx + i);
      if (channels > 1) m *= inv_c;
      acc = fmaf(__ldg(table_t + (long long)k * up + p), m, acc);
    }
    y[j] = acc;
  }
  if (absmax_bits) {
    float a = fabsf(acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) a = fmaxf(a, __shfl_xor_sync(0xffffffffu, a, o));
    __shared__ float red[8];
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int w = 1; w < 8; ++w) a = fmaxf(a, red[w]);
      atomicMax(absmax_bits, __float_as_uint(a));
    }
  }
}

// inference.py:147: audio / max(|audio|.max(), 1)
__global__ void peak_normalize_kernel(float* __restrict__ y, long long n, const unsigned int* __restrict__ absmax_bits,
                                      float floor_) {
  const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const float peak = fmaxf(__uint_as_float(*absmax_bits), floor_);
  y[j] = y[j] / peak;
}
}  // namespace

extern "C" int echo_dwconv_layernorm_ctx(const float* x, int batch, int t, int c, int left_ctx, const float* w_dw,
                                         const float* b_dw, const float* ln_w, const float* ln_b, float eps, void* y,
                                         echo_stream_t stream) {
  ECHO_CHECK_ARG(x && w_dw && b_dw && ln_w && ln_b && y && batch > 0 && t > 0, "dwconv_ln: bad argument");
  ECHO_CHECK_ARG(left_ctx >= 0 && left_ctx <= 6 && (left_ctx == 0 || batch == 1), "dwconv_ln: left_ctx needs batch 1, <= 6 rows");
  ECHO_CHECK_ARG(c > 0 && c <= 1024, "dwconv_ln: c must be <= 1024");
  const int threads = c >= 256 ? 256 : ((c + 31) / 32) * 32;
  dwconv_ln_kernel<<<dim3(t, batch), threads, 0, (cudaStream_t)stream>>>(x, t, c, w_dw, b_dw, ln_w, ln_b, eps, (bf16*)y,
                                                                          left_ctx);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

extern "C" int echo_dwconv_layernorm(const float* x, int batch, int t, int c, const float* w_dw, const float* b_dw,
                                     const float* ln_w, const float* ln_b, float eps, void* y, echo_stream_t stream) {
  return echo_dwconv_layernorm_ctx(x, batch, t, c, 0, w_dw, b_dw, ln_w, ln_b, eps, y, stream);
}

extern "C" int echo_conv_out_tanh_ctx(const void* x, int batch, int t, int c, int left_ctx, const float* w, float bias,
                                      float* wav, echo_stream_t stream) {
  ECHO_CHECK_ARG(x && w && wav && batch > 0 && t > 0 && c > 0 && c <= 128, "conv_out: bad argument");
  ECHO_CHECK_ARG(left_ctx >= 0 && left_ctx <= 6 && (left_ctx == 0 || batch == 1), "conv_out: left_ctx needs batch 1, <= 6 rows");
  const int runs = t >= (1 << 16) ? 4 : 1;                       // 32 * runs outputs per warp
  const int per_cta = 8 * 32 * runs;
  const dim3 grid(ceil_div(t, per_cta), batch);
  cudaStream_t st = (cudaStream_t)stream;
  const bf16* xp = (const bf16*)x;
  ECHO_CHECK_ARG(c % 2 == 0 && ((uintptr_t)x % 4) == 0, "conv_out: channel count must be even, input 4-byte aligned");
  if (c <= 64) conv_out_tanh_kernel<1><<<grid, 256, 0, st>>>(xp, t, c, w, bias, wav, left_ctx, runs);   // pairs across lanes
  else conv_out_tanh_kernel<2><<<grid, 256, 0, st>>>(xp, t, c, w, bias, wav, left_ctx, runs);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

extern "C" int echo_conv_out_tanh(const void* x, int batch, int t, int c, const float* w, float bias, float* wav,
                                  echo_stream_t stream) {
  return echo_conv_out_tanh_ctx(x, batch, t, c, 0, w, bias, wav, stream);
}

extern "C" int echo_cast_snake(const float* x, int64_t rows, int c, const float* alpha, void* y, echo_stream_t stream) {
  ECHO_CHECK_ARG(x && y && rows > 0 && c > 0 && c % 4 == 0, "cast_snake: bad argument");
  const long long n = rows * c;
  cast_snake_kernel<<<(unsigned)((n / 4 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(x, n, c, alpha, (bf16*)y);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

extern "C" int echo_flattening_point(const float* latent, int len, int width, int window, float std_threshold,
                                     float target, int32_t* index_out, echo_stream_t stream) {
  ECHO_CHECK_ARG(latent && index_out && len > 0 && width > 0 && window > 0, "flattening: bad argument");
  const size_t smem = (size_t)(len + window) * 2 * sizeof(double);
  ECHO_CHECK_ARG(smem <= 227 * 1024, "flattening: latent of %d frames too long (prefix sums of %zu bytes exceed 227 KB of shared memory)",
                 len, smem);
  if (smem > 48 * 1024)   // the reference accepts any length (inference.py:315-330); per device, so set on every long call
    ECHO_CUDA(cudaFuncSetAttribute(flattening_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  flattening_kernel<<<1, 256, smem, (cudaStream_t)stream>>>(latent, len, width, window, std_threshold, target, index_out);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

extern "C" int echo_conv_in_snake(const float* audio, int batch, int t, int c, const float* w, const float* bias,
                                  const float* alpha, float* x, void* s, echo_stream_t stream) {
  ECHO_CHECK_ARG(audio && w && bias && alpha && x && s && batch > 0 && t > 0 && c > 0 && c % 4 == 0, "conv_in: bad argument");
  const long long n = (long long)t * (c / 4);
  conv_in_snake_kernel<<<dim3((unsigned)((n + 255) / 256), batch), 256, 0, (cudaStream_t)stream>>>(audio, t, c, w, bias, alpha,
                                                                                                 x, (bf16*)s);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

extern "C" int echo_snake_bf16(const void* x, int64_t rows, int c, const float* alpha, void* y, echo_stream_t stream) {
  ECHO_CHECK_ARG(x && y && alpha && rows > 0 && c > 0 && c % 4 == 0, "snake_bf16: bad argument");
  const long long n = rows * c;
  snake_bf16_kernel<<<(unsigned)((n / 4 + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)x, n, c, alpha, (bf16*)y);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

extern "C" int echo_vq_nearest(const float* e, int m, int dim, const float* codebook_normalized, const float* codebook,
                               int n, int32_t* index_out, float* code_out, echo_stream_t stream) {
  ECHO_CHECK_ARG(e && codebook_normalized && codebook && index_out && m > 0 && n > 0, "vq_nearest: bad argument");
  ECHO_CHECK_ARG(dim == 8, "vq_nearest: codebook_dim %d not supported (8)", dim);
  vq_nearest_kernel<8><<<ceil_div(m, 4), 128, 0, (cudaStream_t)stream>>>(e, m, codebook_normalized, codebook, n, index_out,
                                                                       code_out);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

extern "C" int echo_resample_mono(const float* x, int channels, int64_t n, int64_t ldx, int down, int up, int width,
                                  const float* table_t, const int32_t* tap_lo, const int32_t* tap_hi, float* y,
                                  int64_t n_out, uint32_t* absmax_bits, echo_stream_t stream) {
  ECHO_CHECK_ARG(x && table_t && y && channels > 0 && channels <= 64 && n > 0 && n_out > 0, "resample: bad argument");
  ECHO_CHECK_ARG(down > 0 && up > 0 && width >= 0 && ldx >= n, "resample: bad rates");
  ECHO_CHECK_ARG((tap_lo == nullptr) == (tap_hi == nullptr), "resample: tap_lo and tap_hi go together");
  ECHO_CHECK_ARG(n_out <= ((n + down - 1) / down + 1) * (int64_t)up, "resample: n_out exceeds the frames the input covers");
  const int taps = 2 * width + down;
  const unsigned blocks = (unsigned)((n_out + 255) / 256);
  resample_mono_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, channels, n, ldx, down, up, width, taps, table_t, tap_lo,
                                                                tap_hi, y, n_out, absmax_bits);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

extern "C" int echo_peak_normalize(float* y, int64_t n, const uint32_t* absmax_bits, float floor_, echo_stream_t stream) {
  ECHO_CHECK_ARG(y && absmax_bits && n > 0, "peak_normalize: bad argument");
  peak_normalize_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(y, n, absmax_bits, floor_);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}
