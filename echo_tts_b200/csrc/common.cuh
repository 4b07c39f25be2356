// Shared helpers for the echo_tts_b200 sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#define ECHO_OK 0
#define ECHO_ERR_INVALID (-1)
#define ECHO_ERR_UNSUPPORTED (-2)
#define ECHO_ERR_CUDA (-3)

void echo_set_error(const char* fmt, ...);

#define ECHO_CHECK_ARG(cond, ...)                      \
  do {                                                 \
    if (!(cond)) {                                     \
      echo_set_error(__VA_ARGS__);                     \
      return ECHO_ERR_INVALID;                         \
    }                                                  \
  } while (0)

#define ECHO_CUDA(call)                                                                  \
  do {                                                                                   \
    cudaError_t e__ = (call);                                                            \
    if (e__ != cudaSuccess) {                                                            \
      echo_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
      return ECHO_ERR_CUDA;                                                              \
    }                                                                                    \
  } while (0)

#define ECHO_LAUNCH_CHECK() ECHO_CUDA(cudaPeekAtLastError())

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// launch counter (bench.py reports it as gpu_launches)
extern unsigned long long g_echo_launches;
#define ECHO_COUNT_LAUNCH() (++g_echo_launches)

// ---- programmatic dependent launch (PDL) ----------------------------------------------------------------
// Hot-path kernels are launched with cudaLaunchAttributeProgrammaticStreamSerialization: the next kernel of the
// stream (or CUDA graph) may start its prologue (barrier init, TMEM alloc, tensor-map prefetch) while this one
// drains; it must call pdl_wait() before its first dependent global access.  ECHO_PDL=0 switches it off (A/B).
int echo_pdl_enabled();
#ifdef __CUDACC__
#include <utility>
template <typename... KArgs, typename... Args>
static inline cudaError_t echo_launch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                      Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = echo_pdl_enabled();
  cfg.attrs = attr; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif

typedef __nv_bfloat16 bf16;

__device__ __forceinline__ float bf2f(bf16 v) { return __bfloat162float(v); }
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }
__device__ __forceinline__ float siluf_(float x) { return x / (1.f + __expf(-x)); }
__device__ __forceinline__ float geluf_(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float snakef_(float x, float alpha) {
  float s = sinf(alpha * x);
  return x + s * s / (alpha + 1e-9f);
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
