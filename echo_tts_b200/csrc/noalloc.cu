// Allocation-free split-K GEMM (no counterpart in the reference).
//
// Some callers cannot call cudaMalloc: code under CUDA-graph capture, tight serving loops, harnesses that forbid it.  The
// split-K form of echo_gemm_bf16 needs scratch -- the fp32 partial products [k_splits][rows][n] that the caller normally
// lends (echo_gemm_desc.out_f32 + split_stride) and a consumer kernel sums.  Here that scratch is a file-scope
// `static __device__` array of ECHO_NOALLOC_SCRATCH_BYTES (32 MiB): every GPU that loads the library holds it until the
// process exits, nothing is allocated at call time (no cudaMalloc / cudaFree / allocating library routine), and requests
// whose partial products do not fit are refused with a clear error.
//
// In host code the array's NAME is a host-side stand-in, not a device address; the device address is taken once per
// device with cudaGetSymbolAddress (that device current) and cached.  Every call on a device shares the one array and it
// keeps the last call's contents: nothing in it has to start at zero (each call overwrites every element it later reads),
// but two calls that OVERLAP on the GPU (different streams, or graph branches without an edge between them) race --
// serialise them.
#include "common.cuh"
#include "../../include/echo_b200.h"

namespace {
constexpr long long SCRATCH_BYTES = ECHO_NOALLOC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_noalloc_scratch[SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
float* g_scratch_ptr[MAX_DEVICES] = {};     // device address of g_noalloc_scratch, per device (set-up once, host side)

int scratch_address(float** out) {
  int dev = -1;
  ECHO_CUDA(cudaGetDevice(&dev));
  ECHO_CHECK_ARG(dev >= 0 && dev < MAX_DEVICES, "noalloc: device ordinal %d outside [0, %d)", dev, MAX_DEVICES);
  if (!g_scratch_ptr[dev]) {
    void* p = nullptr;
    ECHO_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    ECHO_CHECK_ARG(p != nullptr, "noalloc: cudaGetSymbolAddress returned a null address");
    g_scratch_ptr[dev] = static_cast<float*>(p);
  }
  *out = g_scratch_ptr[dev];
  return ECHO_OK;
}

// out[m, n] = (resid ? resid[m, n] : 0) + sum_s part[s][m][n]   (fixed summation order: deterministic)
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const float* __restrict__ part, int splits, long long rows, int n,
                                                            const float* resid, long long ld_resid, float* out,
                                                            long long ld_out) {
  pdl_launch_dependents();
  pdl_wait();                                  // the partial products come from the GEMM launched just before
  const int n4 = n >> 2;
  const long long total = rows * n4, stride = rows * (long long)n;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long m = i / n4;
    const int c = (int)(i - m * n4) << 2;
    float4 acc = resid ? *reinterpret_cast<const float4*>(resid + m * ld_resid + c) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float* p = part + m * n + c;
    for (int s = 0; s < splits; ++s) {
      const float4 v = *reinterpret_cast<const float4*>(p + s * stride);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    *reinterpret_cast<float4*>(out + m * ld_out + c) = acc;
  }
}
}  // namespace

extern "C" long long echo_noalloc_scratch_bytes(void) { return SCRATCH_BYTES; }

extern "C" int echo_gemm_bf16_splitk_noalloc(const echo_gemm_desc* d, echo_stream_t stream) {
  ECHO_CHECK_ARG(d != nullptr, "gemm_splitk_noalloc: null descriptor");
  ECHO_CHECK_ARG(d->k_splits >= 2, "gemm_splitk_noalloc: k_splits must be >= 2 (got %d); use echo_gemm_bf16 otherwise", d->k_splits);
  ECHO_CHECK_ARG(d->epilogue == ECHO_EPI_GENERIC && d->act == ECHO_ACT_NONE && !d->out_bf16 && !d->out2_bf16,
                 "gemm_splitk_noalloc: GENERIC epilogue, no activation, fp32 output only");
  ECHO_CHECK_ARG(d->out_f32 && d->ld_out_f32 % 4 == 0 && ((uintptr_t)d->out_f32 % 16) == 0,
                 "gemm_splitk_noalloc: out_f32 missing or not 16-byte aligned (ld %% 4)");
  ECHO_CHECK_ARG(!d->resid || (d->ld_resid % 4 == 0 && ((uintptr_t)d->resid % 16) == 0),
                 "gemm_splitk_noalloc: resid must be 16-byte aligned (ld %% 4)");
  ECHO_CHECK_ARG(d->rows > 0 && d->batch > 0 && d->n > 0 && d->n % 8 == 0, "gemm_splitk_noalloc: bad shape");
  const long long rows = (long long)d->rows * d->batch;
  const long long need = (long long)d->k_splits * rows * d->n * (long long)sizeof(float);
  ECHO_CHECK_ARG(need <= SCRATCH_BYTES,
                 "gemm_splitk_noalloc: k_splits * rows * n * 4 = %lld bytes of partial products exceed the static scratch of "
                 "%lld bytes (ECHO_NOALLOC_SCRATCH_BYTES); use echo_gemm_bf16 with a caller-lent buffer", need, SCRATCH_BYTES);
  float* scratch = nullptr;
  if (int rc = scratch_address(&scratch)) return rc;
  echo_gemm_desc g = *d;                       // the split-K GEMM as it exists, pointed at the static scratch
  g.out_f32 = scratch;
  g.ld_out_f32 = d->n;
  g.split_stride = rows * d->n;
  g.resid = nullptr; g.ld_resid = 0;
  g.workspace = nullptr; g.workspace_bytes = 0;
  if (int rc = echo_gemm_bf16(&g, stream)) return rc;
  const long long total = rows * (d->n / 4);
  const int grid = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  ECHO_CUDA(echo_launch(splitk_reduce_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream, (const float*)scratch,
                        (int)d->k_splits, rows, (int)d->n, (const float*)d->resid, (long long)d->ld_resid, (float*)d->out_f32,
                        (long long)d->ld_out_f32));
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}
