// Error plumbing + the small fp32 / bandwidth-bound kernels of the path.
#include "common.cuh"
#include "../../include/echo_b200.h"
#include <stdarg.h>
#include <stdlib.h>

static thread_local char g_err[512] = "";
unsigned long long g_echo_launches = 0;

void echo_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* echo_last_error(void) { return g_err; }
int echo_pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("ECHO_PDL");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v;
}
extern "C" int echo_version(void) { return 100; }
extern "C" unsigned long long echo_launch_count(void) { return g_echo_launches; }
extern "C" void echo_reset_launch_count(void) { g_echo_launches = 0; }

// ---------------------------------------------------------------------------------------------
// RMS-normalise (+ AdaLN modulate) fp32 rows -> bf16.  One warp per row, float4 loads, row kept
// in registers (d <= 4096).  HBM-bound: 4 B in + 2 B out per element.
// ---------------------------------------------------------------------------------------------
// Four warps per row, two rows per CTA: every load a row needs (x, up to four split-K partials, scale, shift) is in
// flight before the first use -- one memory round trip per row, and 4x the CTAs of a warp-per-row layout when only
// a few hundred rows exist (the single-branch Euler steps).
constexpr int NORM_MAX_DELTA = 4;    // partials in flight at a time (registers); up to 16 are summed
template <int VEC>
__global__ void __launch_bounds__(256) rmsnorm_mod_kernel(const float* __restrict__ x, long long ldx, int m, int d,
                                                          float eps, const float* __restrict__ delta, int n_delta,
                                                          long long delta_stride, float* __restrict__ x_out,
                                                          const float* __restrict__ scale1,
                                                          const float* __restrict__ shift, long long mod_ld,
                                                          int rows_per_mod, const int* __restrict__ step,
                                                          long long mod_step_stride, bf16* __restrict__ y,
                                                          long long ldy) {
  __shared__ float s_part[8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 2 + (warp >> 2);
  const int t = (warp & 3) * 32 + lane;           // 0..127 inside the row
  const int nvec = d >> 2;
  const bool live = row < m;
  pdl_launch_dependents();
  float4 v[VEC], dl[NORM_MAX_DELTA][VEC], sc[VEC], sh[VEC];
  if (live && !step) {   // the modulation table is not produced by the previous kernel: request it before waiting
    const long long g = row / rows_per_mod;
    const float4* scp = reinterpret_cast<const float4*>(scale1 + g * mod_ld);
    const float4* shp = shift ? reinterpret_cast<const float4*>(shift + g * mod_ld) : nullptr;
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
      const int c = t + i * 128;
      if (c < nvec) {
        sc[i] = __ldg(scp + c);
        sh[i] = shp ? __ldg(shp + c) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
  pdl_wait();
  if (live) {
    const float4* xr = reinterpret_cast<const float4*>(x + (long long)row * ldx);
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
      const int c = t + i * 128;
      if (c < nvec) v[i] = xr[c];
    }
#pragma unroll
    for (int s = 0; s < NORM_MAX_DELTA; ++s) {
      if (s < n_delta) {
        const float4* dr = reinterpret_cast<const float4*>(delta + s * delta_stride + (long long)row * ldx);
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
          const int c = t + i * 128;
          if (c < nvec) dl[s][i] = dr[c];
        }
      }
    }
    if (step) {            // table row depends on a device-side step counter the previous kernel may have advanced
      const long long g = row / rows_per_mod + (long long)step[0] * mod_step_stride;
      const float4* scp = reinterpret_cast<const float4*>(scale1 + g * mod_ld);
      const float4* shp = shift ? reinterpret_cast<const float4*>(shift + g * mod_ld) : nullptr;
#pragma unroll
      for (int i = 0; i < VEC; ++i) {
        const int c = t + i * 128;
        if (c < nvec) {
          sc[i] = __ldg(scp + c);
          sh[i] = shp ? __ldg(shp + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
    }
  }
  float ssq = 0.f;
  if (live) {
    if (n_delta > 0) {   // deferred residual update: x += sum of the split-K partial outputs (fixed order), written back
#pragma unroll
      for (int s = 0; s < NORM_MAX_DELTA; ++s) {
        if (s < n_delta) {
#pragma unroll
          for (int i = 0; i < VEC; ++i) {
            if (t + i * 128 < nvec) { v[i].x += dl[s][i].x; v[i].y += dl[s][i].y; v[i].z += dl[s][i].z; v[i].w += dl[s][i].w; }
          }
        }
      }
      for (int s0 = NORM_MAX_DELTA; s0 < n_delta; s0 += NORM_MAX_DELTA) {   // more partials: four at a time, same order
#pragma unroll
        for (int s = 0; s < NORM_MAX_DELTA; ++s) {
          if (s0 + s < n_delta) {
            const float4* dr = reinterpret_cast<const float4*>(delta + (s0 + s) * delta_stride + (long long)row * ldx);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
              const int c = t + i * 128;
              if (c < nvec) dl[s][i] = dr[c];
            }
          }
        }
#pragma unroll
        for (int s = 0; s < NORM_MAX_DELTA; ++s) {
          if (s0 + s < n_delta) {
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
              if (t + i * 128 < nvec) { v[i].x += dl[s][i].x; v[i].y += dl[s][i].y; v[i].z += dl[s][i].z; v[i].w += dl[s][i].w; }
            }
          }
        }
      }
      float4* xo = reinterpret_cast<float4*>(x_out + (long long)row * ldx);
#pragma unroll
      for (int i = 0; i < VEC; ++i) {
        const int c = t + i * 128;
        if (c < nvec) xo[c] = v[i];
      }
    }
#pragma unroll
    for (int i = 0; i < VEC; ++i)
      if (t + i * 128 < nvec) ssq += v[i].x * v[i].x + v[i].y * v[i].y + v[i].z * v[i].z + v[i].w * v[i].w;
  }
  ssq = warp_sum(ssq);
  if (lane == 0) s_part[warp] = ssq;
  __syncthreads();
  if (!live) return;
  const int w0 = warp & ~3;
  const float tot = (s_part[w0] + s_part[w0 + 1]) + (s_part[w0 + 2] + s_part[w0 + 3]);
  const float rstd = rsqrtf(tot / (float)d + eps);
  uint2* yr = reinterpret_cast<uint2*>(y + (long long)row * ldy);
#pragma unroll
  for (int i = 0; i < VEC; ++i) {
    const int c = t + i * 128;
    if (c < nvec) {
      const float4 o = make_float4(fmaf(v[i].x * rstd, sc[i].x, sh[i].x), fmaf(v[i].y * rstd, sc[i].y, sh[i].y),
                                   fmaf(v[i].z * rstd, sc[i].z, sh[i].z), fmaf(v[i].w * rstd, sc[i].w, sh[i].w));
      yr[c] = make_uint2(pack_bf16x2(o.x, o.y), pack_bf16x2(o.z, o.w));
    }
  }
}

static int launch_rmsnorm(const float* x, int64_t ldx, int m, int d, float eps, const float* delta, int n_delta,
                          int64_t delta_stride, float* x_out, const float* scale1, const float* shift, int64_t mod_ld,
                          int rows_per_mod, const int32_t* step, int64_t mod_step_stride, void* y, int64_t ldy,
                          echo_stream_t stream) {
  ECHO_CHECK_ARG(x && scale1 && y && m > 0, "rmsnorm: null argument");
  ECHO_CHECK_ARG(n_delta >= 0 && n_delta <= 16 && (n_delta == 0 || (delta && x_out && delta_stride % 4 == 0)),
                 "rmsnorm: bad delta arguments (at most 16 partials)");
  ECHO_CHECK_ARG(d % 4 == 0 && d <= 4096 && ldx % 4 == 0 && ldy % 4 == 0 && mod_ld % 4 == 0, "rmsnorm: d must be <=4096, %%4");
  if (rows_per_mod <= 0) rows_per_mod = 1 << 30;
  dim3 grid(ceil_div(m, 2));
  cudaStream_t st = (cudaStream_t)stream;
  if (d <= 1024)
    ECHO_CUDA(echo_launch(rmsnorm_mod_kernel<2>, grid, dim3(256), 0, st, x, (long long)ldx, m, d, eps, delta, n_delta, (long long)delta_stride, x_out, scale1, shift, (long long)mod_ld, rows_per_mod, step, (long long)mod_step_stride, (bf16*)y, (long long)ldy));
  else if (d <= 2048)
    ECHO_CUDA(echo_launch(rmsnorm_mod_kernel<4>, grid, dim3(256), 0, st, x, (long long)ldx, m, d, eps, delta, n_delta, (long long)delta_stride, x_out, scale1, shift, (long long)mod_ld, rows_per_mod, step, (long long)mod_step_stride, (bf16*)y, (long long)ldy));
  else
    ECHO_CUDA(echo_launch(rmsnorm_mod_kernel<8>, grid, dim3(256), 0, st, x, (long long)ldx, m, d, eps, delta, n_delta, (long long)delta_stride, x_out, scale1, shift, (long long)mod_ld, rows_per_mod, step, (long long)mod_step_stride, (bf16*)y, (long long)ldy));
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

extern "C" int echo_rmsnorm_mod(const float* x, int64_t ldx, int m, int d, float eps, const float* scale1,
                                const float* shift, int64_t mod_ld, int rows_per_mod, const int32_t* step,
                                int64_t mod_step_stride, void* y, int64_t ldy, echo_stream_t stream) {
  return launch_rmsnorm(x, ldx, m, d, eps, nullptr, 0, 0, nullptr, scale1, shift, mod_ld, rows_per_mod, step, mod_step_stride,
                        y, ldy, stream);
}
extern "C" int echo_add_rmsnorm_mod(float* x, int64_t ldx, int m, int d, float eps, const float* delta, int n_delta,
                                    int64_t delta_stride, const float* scale1, const float* shift, int64_t mod_ld,
                                    int rows_per_mod, void* y, int64_t ldy, echo_stream_t stream) {
  return launch_rmsnorm(x, ldx, m, d, eps, delta, n_delta, delta_stride, x, scale1, shift, mod_ld, rows_per_mod, nullptr, 0, y,
                        ldy, stream);
}

// RMSNorm with an fp32 result (autoencoder.py:726-731): the quantizer's input must not be rounded to bf16 (the
// nearest-code search is precision-sensitive).  One warp per row.
__global__ void rmsnorm_f32_kernel(const float* __restrict__ x, long long ldx, int m, int d, float eps,
                                   const float* __restrict__ w, float* __restrict__ y, long long ldy) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= m) return;
  const float* xr = x + (long long)row * ldx;
  float ssq = 0.f;
  for (int c = lane; c < d; c += 32) ssq = fmaf(xr[c], xr[c], ssq);
  const float rstd = rsqrtf(warp_sum(ssq) / (float)d + eps);
  float* yr = y + (long long)row * ldy;
  for (int c = lane; c < d; c += 32) yr[c] = xr[c] * rstd * w[c];
}
extern "C" int echo_rmsnorm_f32(const float* x, int64_t ldx, int m, int d, float eps, const float* weight, float* y,
                                int64_t ldy, echo_stream_t stream) {
  ECHO_CHECK_ARG(x && weight && y && m > 0 && d > 0, "rmsnorm_f32: bad argument");
  rmsnorm_f32_kernel<<<ceil_div(m, 8), 256, 0, (cudaStream_t)stream>>>(x, (long long)ldx, m, d, eps, weight, y, (long long)ldy);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

// ---------------------------------------------------------------------------------------------
// fp32 CUDA-core GEMM: 64x64 tile, 16x16 threads, 4x4 micro-tile.  Small / accuracy-critical products; the big ones
// are weight streams with a handful of rows (the AdaLN modulation table: 40 timesteps x 600 MB of fp32 weights per
// generation), so the loop is built for memory-level parallelism: K step 32, 16-byte loads, the next tile requested into
// registers before the current one is multiplied (VEC; K, the leading dimensions and the bases must be multiples of 4
// floats -- otherwise the scalar K-step-16 form).  Measured on the table's down-projection (40 x 12288 x 2048): 400 -> see
// DESIGN.md.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void sgemm_store(const echo_sgemm_desc& D, int z, int m0, int n0, int tx, int ty, float (&acc)[4][4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= D.m) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= D.n) continue;
      float v = acc[i][j];
      if (D.bias) v += D.bias[z * D.bias_zstride + n];
      v *= D.out_scale;
      if (D.addend) v += D.addend[z * D.addend_zstride + (long long)m * D.ld_addend + n];
      if (D.post == ECHO_POST_PLUS1) v += 1.f;
      else if (D.post == ECHO_POST_TANH) v = tanhf(v);
      else if (D.post == ECHO_POST_SILU) v = v / (1.f + expf(-v));
      if (D.out) D.out[z * D.out_zstride + (long long)m * D.ldo + n] = v;
      if (D.out_bf16) ((bf16*)D.out_bf16)[z * D.out_bf16_zstride + (long long)m * D.ldo_bf16 + n] = __float2bfloat16(v);
    }
  }
}

template <bool VEC>
__global__ void __launch_bounds__(256) sgemm_kernel(const __grid_constant__ echo_sgemm_desc D) {
  constexpr int KS = VEC ? 32 : 16;
  __shared__ float sA[KS][64 + 4];
  __shared__ float sW[KS][64 + 4];
  const int z = blockIdx.z;
  // NO early trigger here (the implicit one at grid exit applies): this kernel writes the AdaLN modulation table, which
  // rmsnorm_mod_kernel requests BEFORE its own griddepcontrol.wait.  With an early trigger, a kernel two launches behind
  // (table sgemm -> in_proj sgemm -> norm, the model.__call__ path) could be resident while the table is still being
  // written; without it every kernel launched after an sgemm starts only once that sgemm has completed.
  pdl_wait();
  const float* A = D.a + z * D.a_zstride;
  const float* W = D.w + z * D.w_zstride;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4] = {};
  if constexpr (VEC) {
    // thread -> two (row, 4-float k group) cells of each operand tile: rows r0 and r0 + 32, k group kq
    const int r0 = threadIdx.x >> 3, kq = (threadIdx.x & 7) * 4;
    float4 ra[2], rw[2];
    auto fetch = [&](int k0) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int r = r0 + 32 * h, k = k0 + kq;
        ra[h] = (k < D.k && m0 + r < D.m) ? __ldg(reinterpret_cast<const float4*>(A + (long long)(m0 + r) * D.lda + k))
                                          : make_float4(0.f, 0.f, 0.f, 0.f);
        rw[h] = (k < D.k && n0 + r < D.n) ? __ldg(reinterpret_cast<const float4*>(W + (long long)(n0 + r) * D.ldw + k))
                                          : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    };
    fetch(0);
    for (int k0 = 0; k0 < D.k; k0 += KS) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int r = r0 + 32 * h;
        float4 a = ra[h];
        if (D.pre_silu) {
          a.x = a.x / (1.f + expf(-a.x)); a.y = a.y / (1.f + expf(-a.y));
          a.z = a.z / (1.f + expf(-a.z)); a.w = a.w / (1.f + expf(-a.w));
        }
        sA[kq][r] = a.x; sA[kq + 1][r] = a.y; sA[kq + 2][r] = a.z; sA[kq + 3][r] = a.w;
        sW[kq][r] = rw[h].x; sW[kq + 1][r] = rw[h].y; sW[kq + 2][r] = rw[h].z; sW[kq + 3][r] = rw[h].w;
      }
      __syncthreads();
      if (k0 + KS < D.k) fetch(k0 + KS);          // in flight under the multiplies below
#pragma unroll
      for (int kk = 0; kk < KS; ++kk) {
        float a[4], w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { a[i] = sA[kk][ty * 4 + i]; w[i] = sW[kk][tx * 4 + i]; }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
      }
      __syncthreads();
    }
  } else {
    for (int k0 = 0; k0 < D.k; k0 += KS) {
      for (int i = threadIdx.x; i < 64 * KS; i += 256) {
        const int r = i / KS, kk = i % KS;
        const int k = k0 + kk;
        float a = 0.f, w = 0.f;
        if (k < D.k) {
          if (m0 + r < D.m) {
            a = A[(long long)(m0 + r) * D.lda + k];
            if (D.pre_silu) a = a / (1.f + expf(-a));
          }
          if (n0 + r < D.n) w = W[(long long)(n0 + r) * D.ldw + k];
        }
        sA[kk][r] = a;
        sW[kk][r] = w;
      }
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < KS; ++kk) {
        float a[4], w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { a[i] = sA[kk][ty * 4 + i]; w[i] = sW[kk][tx * 4 + i]; }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
      }
      __syncthreads();
    }
  }
  sgemm_store(D, z, m0, n0, tx, ty, acc);
}

extern "C" int echo_sgemm_f32(const echo_sgemm_desc* d, echo_stream_t stream) {
  ECHO_CHECK_ARG(d && d->a && d->w && (d->out || d->out_bf16), "sgemm: null argument");
  ECHO_CHECK_ARG(d->m > 0 && d->n > 0 && d->k > 0 && d->nz > 0, "sgemm: bad shape");
  dim3 grid(ceil_div(d->n, 64), ceil_div(d->m, 64), d->nz);
  const bool vec = d->k % 4 == 0 && d->lda % 4 == 0 && d->ldw % 4 == 0 && d->a_zstride % 4 == 0 && d->w_zstride % 4 == 0 &&
                   ((uintptr_t)d->a % 16) == 0 && ((uintptr_t)d->w % 16) == 0;
  if (vec) ECHO_CUDA(echo_launch(sgemm_kernel<true>, grid, dim3(256), 0, (cudaStream_t)stream, *d));
  else ECHO_CUDA(echo_launch(sgemm_kernel<false>, grid, dim3(256), 0, (cudaStream_t)stream, *d));
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

// timestep embedding, model.py:27-43 (frequencies computed exactly like torch: exp(-ln(1e4) * j / half) * 1000)
__global__ void timestep_embedding_kernel(const float* t, int n, int size, float* out) {
  const int half = size / 2;
  const int i = blockIdx.x, j = threadIdx.x + blockIdx.y * blockDim.x;
  if (j >= half) return;
  const float w = 1000.f * expf(-logf(10000.f) * (float)j / (float)half);
  const float a = t[i] * w;
  out[(long long)i * size + j] = cosf(a);
  out[(long long)i * size + half + j] = sinf(a);
}
extern "C" int echo_timestep_embedding(const float* t, int n, int size, float* out, echo_stream_t stream) {
  ECHO_CHECK_ARG(t && out && n > 0 && size > 0 && size % 2 == 0, "timestep_embedding: bad argument");
  dim3 grid(n, ceil_div(size / 2, 128));
  timestep_embedding_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(t, n, size, out);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

__global__ void embedding_kernel(const int* ids, int n, const float* table, int d, float* out) {
  const int i = blockIdx.x;
  const float4* src = reinterpret_cast<const float4*>(table + (long long)ids[i] * d);
  float4* dst = reinterpret_cast<float4*>(out + (long long)i * d);
  for (int c = threadIdx.x; c < d / 4; c += blockDim.x) dst[c] = src[c];
}
extern "C" int echo_embedding(const int32_t* ids, int n, const float* table, int d, float* out, echo_stream_t stream) {
  ECHO_CHECK_ARG(ids && table && out && n > 0 && d % 4 == 0, "embedding: bad argument");
  embedding_kernel<<<n, 128, 0, (cudaStream_t)stream>>>(ids, n, table, d, out);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

__global__ void mask_extent_kernel(const uint8_t* mask, int len, long long stride, int* extent) {
  const uint8_t* row = mask + (long long)blockIdx.x * stride;
  int best = 0;
  for (int j = threadIdx.x; j < len; j += blockDim.x)
    if (row[j]) best = max(best, j + 1);
  __shared__ int s[32];
  for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = best;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) best = max(best, s[w]);
    extent[blockIdx.x] = best;
  }
}
extern "C" int echo_mask_extent(const uint8_t* mask, int rows, int len, int64_t stride, int32_t* extent,
                                echo_stream_t stream) {
  ECHO_CHECK_ARG(mask && extent && rows > 0 && len >= 0, "mask_extent: bad argument");
  mask_extent_kernel<<<rows, 256, 0, (cudaStream_t)stream>>>(mask, len, stride, extent);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

__global__ void step_advance_kernel(int* step) { step[0] += 1; }
extern "C" int echo_step_advance(int32_t* step, echo_stream_t stream) {
  ECHO_CHECK_ARG(step, "step_advance: null");
  step_advance_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}
