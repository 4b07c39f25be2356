// Fused sampler tail (fp32): out_norm -> out_proj -> CFG combine -> temporal score rescale -> Euler.
// One CTA owns PP latent positions x all CFG branches: the normalised rows live in shared memory; each warp streams
// CW out_proj weight rows at a time (L2-resident, 80 x d fp32) against every row -- a [BR*PP x CW] register tile per
// lane, lanes splitting K, so one weight float4 and one row float4 feed CW and BR*PP FMAs x 4 respectively -- and
// reduces with warp shuffles; the combine/update happens in registers.  128 CTAs for 640 positions: one wave.
#include "common.cuh"
#include "../../include/echo_b200.h"

namespace {
constexpr int PP = 5, CW = 5;

// v = v_c + a (v_c - v_ut) + b (v_c - v_us)  (inference.py:526-530), optional temporal score rescale (inference.py:431-443),
// x += v dt (inference.py:558)
// guidance constants: immediates of the descriptor, or (graph replay with per-request values) a device vector
struct Guidance { float cfg_text, cfg_speaker, rescale_k, rescale_sigma; };
__device__ __forceinline__ Guidance load_guidance(const echo_tail_desc& D) {
  if (D.guidance) return Guidance{D.guidance[0], D.guidance[1], D.guidance[2], D.guidance[3]};
  return Guidance{D.cfg_text, D.cfg_speaker, D.rescale_k, D.rescale_sigma};
}
__device__ __forceinline__ float combine3(const Guidance& G, float vc, float vt, float vs) {
  return vc + G.cfg_text * (vc - vt) + G.cfg_speaker * (vc - vs);
}
__device__ __forceinline__ float rescale_and_step(const echo_tail_desc& D, const Guidance& G, float v, float t, float dt,
                                                  long long xi) {
  if (D.update_x) {
    const float x = D.x_t[xi];
    if (D.use_rescale && t < 1.f) {
      const float omt = 1.f - t;
      const float snr = (omt * omt) / (t * t);
      const float sg2 = G.rescale_sigma * G.rescale_sigma;
      const float ratio = (snr * sg2 + 1.f) / (snr * sg2 / G.rescale_k + 1.f);
      v = 1.f / omt * (ratio * (omt * v + x) - x);
    }
    D.x_t[xi] = x + v * dt;
  }
  if (D.v_comb_out) D.v_comb_out[xi] = v;
  return v;
}

// Combine-only variant for the multi-GPU CFG-branch split: v_out holds the three gathered branch velocities
// fp32 [3][positions][latent] (INPUT); everything else as in the fused tail.
__global__ void __launch_bounds__(256) cfg_combine_euler_kernel(const __grid_constant__ echo_tail_desc D) {
  pdl_launch_dependents();
  pdl_wait();
  if (D.wait_flag[0]) {
    // fused exchange: the other two ranks' tail kernels pushed their branch velocities into our v_out over NVLink and
    // then bumped these counters (release, system scope).  Bounded wait (~1 minute: a peer that is merely late -- first-call
    // warm-up, a paused rank -- is waited for; a lost one traps instead of hanging the GPU for good).
    if (threadIdx.x < 2) {
      const int32_t* f = D.wait_flag[threadIdx.x];
      int v = 0;
      const long long t0 = clock64();
      do {
        asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
        if (v < D.wait_value && clock64() - t0 > 120000000000LL) { printf("echo: CFG-split peer never arrived\n"); __trap(); }
      } while (v < D.wait_value);
    }
    __syncthreads();
  }
  const int s = D.step ? D.step[0] : 0;
  const float t = D.sched[s * 4 + 0], dt = D.sched[s * 4 + 1];
  const long long n = (long long)D.positions * D.latent;
  const Guidance G = load_guidance(D);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    rescale_and_step(D, G, combine3(G, D.v_out[i], D.v_out[n + i], D.v_out[2 * n + i]), t, dt, i);
}

template <int BR>
__global__ void __launch_bounds__(256) cfg_euler_tail_kernel(const __grid_constant__ echo_tail_desc D) {
  extern __shared__ float sm[];
  const int nrows = D.branches * PP;
  float* srow = sm;                       // [nrows][d]
  float* sv = sm + (size_t)nrows * D.d;   // [nrows][latent]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int pos0 = blockIdx.x * PP;
  const int nvec = D.d >> 2;
  pdl_launch_dependents();
  pdl_wait();

  for (int r = warp; r < nrows; r += 8) {
    const int br = r / PP, p = pos0 + (r % PP);
    float4* dst = reinterpret_cast<float4*>(srow + (size_t)r * D.d);
    if (p >= D.positions) {
      for (int c = lane; c < nvec; c += 32) dst[c] = make_float4(0.f, 0.f, 0.f, 0.f);
      continue;
    }
    const long long roff = ((long long)br * D.positions + p) * D.ldh;
    const float4* src = reinterpret_cast<const float4*>(D.h + roff);
    float ssq = 0.f;
    for (int c = lane; c < nvec; c += 32) {
      float4 v = src[c];
      for (int s = 0; s < D.n_delta; ++s) {   // deferred residual update of the last projection (fixed order)
        const float4 dv = reinterpret_cast<const float4*>(D.delta + s * D.delta_stride + roff)[c];
        v.x += dv.x; v.y += dv.y; v.z += dv.z; v.w += dv.w;
      }
      dst[c] = v;
      ssq += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    ssq = warp_sum(ssq);
    const float rstd = rsqrtf(ssq / (float)D.d + D.eps);
    const float4* nw = reinterpret_cast<const float4*>(D.norm_w);
    for (int c = lane; c < nvec; c += 32) {
      const float4 v = dst[c], w = __ldg(nw + c);
      dst[c] = make_float4(v.x * rstd * w.x, v.y * rstd * w.y, v.z * rstd * w.z, v.w * rstd * w.w);
    }
  }
  __syncthreads();

  for (int o0 = warp * CW; o0 < D.latent; o0 += 8 * CW) {
    float acc[BR * PP][CW];
#pragma unroll
    for (int r = 0; r < BR * PP; ++r)
#pragma unroll
      for (int j = 0; j < CW; ++j) acc[r][j] = 0.f;
    const float4* w = reinterpret_cast<const float4*>(D.w_out + (long long)o0 * D.d);
    for (int c = lane; c < nvec; c += 32) {
      float4 wv[CW];
#pragma unroll
      for (int j = 0; j < CW; ++j) wv[j] = (o0 + j < D.latent) ? __ldg(w + (long long)j * nvec + c) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int r = 0; r < BR * PP; ++r) {
        const float4 xv = reinterpret_cast<const float4*>(srow + (size_t)r * D.d)[c];
#pragma unroll
        for (int j = 0; j < CW; ++j)
          acc[r][j] = fmaf(wv[j].x, xv.x, fmaf(wv[j].y, xv.y, fmaf(wv[j].z, xv.z, fmaf(wv[j].w, xv.w, acc[r][j]))));
      }
    }
#pragma unroll
    for (int j = 0; j < CW; ++j) {
      const float b = (o0 + j < D.latent) ? __ldg(D.b_out + o0 + j) : 0.f;
#pragma unroll
      for (int r = 0; r < BR * PP; ++r) {
        const float sum = warp_sum(acc[r][j]);
        if (lane == 0 && o0 + j < D.latent) sv[r * D.latent + o0 + j] = sum + b;
      }
    }
  }
  __syncthreads();

  const int s = D.step ? D.step[0] : 0;
  const float t = D.sched ? D.sched[s * 4 + 0] : 0.f, dt = D.sched ? D.sched[s * 4 + 1] : 0.f;
  const Guidance G = load_guidance(D);
  for (int i = threadIdx.x; i < PP * D.latent; i += blockDim.x) {
    const int pl = i / D.latent, c = i % D.latent;
    const int p = pos0 + pl;
    if (p >= D.positions) continue;
    const float vc = sv[(0 * PP + pl) * D.latent + c];
    float v = vc;
    if (D.v_out) D.v_out[((long long)0 * D.positions + p) * D.latent + c] = vc;
    if (D.v_peer[0]) {      // branches == 1: push this rank's branch velocity into the two other ranks' gather buffers
      D.v_peer[0][(long long)p * D.latent + c] = vc;
      D.v_peer[1][(long long)p * D.latent + c] = vc;
    }
    if (D.branches == 3) {
      const float vt = sv[(1 * PP + pl) * D.latent + c];
      const float vs = sv[(2 * PP + pl) * D.latent + c];
      if (D.v_out) {
        D.v_out[((long long)1 * D.positions + p) * D.latent + c] = vt;
        D.v_out[((long long)2 * D.positions + p) * D.latent + c] = vs;
      }
      v = combine3(G, vc, vt, vs);
    }
    rescale_and_step(D, G, v, t, dt, (long long)p * D.latent + c);
  }
  if (D.flag_peer[0]) {
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < 2)
      asm volatile("red.release.sys.global.add.s32 [%0], %1;" ::"l"(D.flag_peer[threadIdx.x]), "r"(1) : "memory");
  }
}
}  // namespace

extern "C" int echo_cfg_euler_tail(const echo_tail_desc* d, echo_stream_t stream) {
  ECHO_CHECK_ARG(d && d->h && d->norm_w && d->w_out && d->b_out, "tail: null argument");
  ECHO_CHECK_ARG(d->branches == 1 || d->branches == 3, "tail: branches must be 1 or 3");
  ECHO_CHECK_ARG(d->d % 4 == 0 && d->ldh % 4 == 0 && d->positions > 0 && d->latent > 0, "tail: bad shape");
  ECHO_CHECK_ARG(!d->update_x || (d->x_t && d->sched), "tail: x_t/sched required when update_x");
  ECHO_CHECK_ARG(d->update_x || d->v_out || d->v_comb_out, "tail: nothing to write");
  ECHO_CHECK_ARG(d->n_delta >= 0 && (d->n_delta == 0 || (d->delta && d->delta_stride % 4 == 0)), "tail: bad delta arguments");
  ECHO_CHECK_ARG(!d->v_peer[0] || (d->branches == 1 && d->v_peer[1] && d->flag_peer[0] && d->flag_peer[1]),
                 "tail: the fused exchange needs branches == 1, two peer buffers and two peer counters");
  const size_t smem = ((size_t)d->branches * PP * (d->d + d->latent)) * sizeof(float);
  ECHO_CHECK_ARG(smem <= 220 * 1024, "tail: d = %d too wide for the shared-memory row cache", d->d);
  // the attribute is per device: set it on every launch (a host-side table write, no device work)
  ECHO_CUDA(cudaFuncSetAttribute(cfg_euler_tail_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ECHO_CUDA(cudaFuncSetAttribute(cfg_euler_tail_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (d->branches == 3)
    ECHO_CUDA(echo_launch(cfg_euler_tail_kernel<3>, dim3(ceil_div(d->positions, PP)), dim3(256), smem, (cudaStream_t)stream, *d));
  else
    ECHO_CUDA(echo_launch(cfg_euler_tail_kernel<1>, dim3(ceil_div(d->positions, PP)), dim3(256), smem, (cudaStream_t)stream, *d));
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

extern "C" int echo_cfg_combine_euler(const echo_tail_desc* d, echo_stream_t stream) {
  ECHO_CHECK_ARG(d && d->v_out && d->sched && d->positions > 0 && d->latent > 0, "combine: null argument");
  ECHO_CHECK_ARG(!d->update_x || d->x_t, "combine: x_t required when update_x");
  ECHO_CHECK_ARG(!d->wait_flag[0] || d->wait_flag[1], "combine: two wait counters or none");
  const long long n = (long long)d->positions * d->latent;
  const int grid = (int)((n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184);
  ECHO_CUDA(echo_launch(cfg_combine_euler_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream, *d));
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

/* CTAs echo_cfg_euler_tail launches for `positions` rows (each adds 1 to the peer counters of the fused exchange). */
extern "C" int echo_cfg_tail_ctas(int positions) { return ceil_div(positions, PP); }
