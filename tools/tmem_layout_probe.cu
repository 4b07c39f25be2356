// Probe (development aid, not part of the library): which accumulator element does each thread get from the
// tcgen05.ld shapes other than 32x32b?  Lane r of TMEM, column c is filled with 1000*r + c through tcgen05.st.32x32b
// (row-per-thread layout, the one csrc/ uses), then read back with .16x256b.x1 / .16x128b.x2 / .16x64b.x4 and printed.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/build/tmem_layout_probe tools/tmem_layout_probe.cu && tools/build/tmem_layout_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void probe(float* out) {
  __shared__ uint32_t slot;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = slot + ((uint32_t)(warp * 32) << 16);     // this warp's lane quarter
  // fill: lane r (= warp*32 + lane), columns 0..31
  {
    uint32_t v[32];
    for (int c = 0; c < 32; ++c) v[c] = __float_as_uint(1000.f * (warp * 32 + lane) + c);
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,"
        "%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(base),
        "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
        "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]),
        "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]),
        "r"(v[31])
        : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (warp == 0) {
    uint32_t r[4];
    // 16x256b.x1: 16 lanes x 8 columns, 4 registers per thread
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(base));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int i = 0; i < 4; ++i) out[0 * 128 + lane * 4 + i] = __uint_as_float(r[i]);
    // same shape, second half of the lane quarter (lanes 16..31)
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(base + (16u << 16)));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int i = 0; i < 4; ++i) out[1 * 128 + lane * 4 + i] = __uint_as_float(r[i]);
    // 16x128b.x2: 16 lanes x 8 columns as two 4-column halves
    asm volatile("tcgen05.ld.sync.aligned.16x128b.x2.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(base));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int i = 0; i < 4; ++i) out[2 * 128 + lane * 4 + i] = __uint_as_float(r[i]);
    // 16x64b.x4
    asm volatile("tcgen05.ld.sync.aligned.16x64b.x4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(base));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int i = 0; i < 4; ++i) out[3 * 128 + lane * 4 + i] = __uint_as_float(r[i]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(slot) : "memory");
}

int main() {
  float* d;
  cudaMalloc(&d, 4 * 128 * sizeof(float));
  cudaMemset(d, 0, 4 * 128 * sizeof(float));
  probe<<<1, 128>>>(d);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
  float h[4 * 128];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  const char* names[4] = {"16x256b.x1 @lane0", "16x256b.x1 @lane16", "16x128b.x2 @lane0", "16x64b.x4 @lane0"};
  for (int s = 0; s < 4; ++s) {
    printf("== %s: thread -> (row,col) of registers r0..r3\n", names[s]);
    for (int t = 0; t < 32; ++t) {
      printf("  t%02d:", t);
      for (int i = 0; i < 4; ++i) { int v = (int)h[s * 128 + t * 4 + i]; printf(" (%d,%d)", v / 1000, v % 1000); }
      printf("\n");
    }
  }
  return 0;
}
