// Legacy tensor-core path (mma.sync.m16n8k16 + cp.async + ldmatrix), 128x128x32 tiles, 8 warps.
// Not the product path: kept as the "recompiled pre-Blackwell kernel" baseline the tcgen05 kernel is
// measured against, and as a debugging aid.  Shares the fused epilogues through a shared-memory stage.
#include "common.cuh"
#include "gemm_epilogue.cuh"


namespace {
constexpr int LBM = 128, LBN = 128, LBK = 32;
constexpr int C_LD = LBN + 4;
constexpr int TILE_BYTES = LBM * LBK * 2;  // 8 KB per operand per stage
constexpr int LEG_SMEM = 4 * TILE_BYTES + LBM * C_LD * 4;

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
__device__ __forceinline__ void mma16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// tile row r (64 B = 4 chunks of 16 B); chunk index XOR-swizzled so ldmatrix rows hit distinct banks
__device__ __forceinline__ uint32_t tile_off(int r, int chunk) { return r * 64 + ((chunk ^ ((r >> 1) & 3)) << 4); }

struct SmemLoader {
  const float* row;
  __device__ __forceinline__ void load32(int col, float* v) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = row[col + j];
  }
};

template <int EPI>
__global__ void __launch_bounds__(256) gemm_legacy_kernel(const __grid_constant__ GemmShape G,
                                                          const __grid_constant__ EpiParams P) {
  extern __shared__ __align__(128) uint8_t smem[];
  float* sC = reinterpret_cast<float*>(smem + 4 * TILE_BYTES);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n0 = blockIdx.x * LBN, m0 = blockIdx.y * LBM, batch = blockIdx.z;
  const int wm = (warp >> 2) * 64, wn = (warp & 3) * 32;
  const bf16* A = G.a + (long long)batch * G.a_batch_stride;
  const int kbt = (G.k_tap + LBK - 1) / LBK;
  const int total = kbt * G.taps;
  const uint32_t sbase = smem_u32(smem);

  float acc[4][4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int k = 0; k < 4; ++k) acc[i][j][k] = 0.f;

  auto issue = [&](int kb, int stage) {
    const int tap = kb / kbt, k0 = (kb - tap * kbt) * LBK;
    const uint32_t sa = sbase + stage * 2 * TILE_BYTES, sb = sa + TILE_BYTES;
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int idx = tid + it * 256;       // 512 chunks per operand tile
      const int r = idx >> 2, c = idx & 3;
      const int k = k0 + c * 8;
      const int ar = m0 + r + G.tap_shift[tap];
      const bool av = (ar >= 0) && (ar < G.rows + G.a_ctx) && (k < G.k_tap);
      cp_async16(sa + tile_off(r, c), av ? (const void*)(A + (long long)ar * G.lda + k) : (const void*)G.a, av);
      const int br = n0 + r;
      const bool bv = (br < G.n) && (k < G.k_tap);
      cp_async16(sb + tile_off(r, c), bv ? (const void*)(G.b + (long long)br * G.ldb + tap * G.k_tap + k) : (const void*)G.b, bv);
    }
    cp_async_commit();
  };

  issue(0, 0);
  for (int kb = 0; kb < total; ++kb) {
    const int stage = kb & 1;
    if (kb + 1 < total) { issue(kb + 1, stage ^ 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    const uint32_t sa = sbase + stage * 2 * TILE_BYTES, sb = sa + TILE_BYTES;
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
      uint32_t af[4][4], bfr[4][2];
#pragma unroll
      for (int mi = 0; mi < 4; ++mi) {
        const int r = wm + mi * 16 + (lane & 15), c = ks * 2 + (lane >> 4);
        ldsm_x4(sa + tile_off(r, c), af[mi][0], af[mi][1], af[mi][2], af[mi][3]);
      }
#pragma unroll
      for (int nj = 0; nj < 2; ++nj) {
        const int r = wn + nj * 16 + (lane & 7) + ((lane >> 4) << 3), c = ks * 2 + ((lane >> 3) & 1);
        ldsm_x4(sb + tile_off(r, c), bfr[nj * 2][0], bfr[nj * 2][1], bfr[nj * 2 + 1][0], bfr[nj * 2 + 1][1]);
      }
#pragma unroll
      for (int mi = 0; mi < 4; ++mi)
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) mma16816(acc[mi][ni], af[mi], bfr[ni][0], bfr[ni][1]);
    }
    __syncthreads();
  }
  // accumulators -> shared memory (fp32), then one thread per row runs the shared epilogue
#pragma unroll
  for (int mi = 0; mi < 4; ++mi)
#pragma unroll
    for (int ni = 0; ni < 4; ++ni) {
      const int r = wm + mi * 16 + (lane >> 2), c = wn + ni * 8 + (lane & 3) * 2;
      sC[r * C_LD + c] = acc[mi][ni][0];
      sC[r * C_LD + c + 1] = acc[mi][ni][1];
      sC[(r + 8) * C_LD + c] = acc[mi][ni][2];
      sC[(r + 8) * C_LD + c + 1] = acc[mi][ni][3];
    }
  __syncthreads();
  if (tid < LBM) {
    SmemLoader L{sC + tid * C_LD};
    run_epilogue<EPI>(LBN, P, L, batch, m0 + tid, n0);
  }
}
}  // namespace

template <int EPI>
int launch_legacy(const GemmShape& G, const EpiParams& P, int batch, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    ECHO_CUDA(cudaFuncSetAttribute(gemm_legacy_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, LEG_SMEM));
    attr_set = true;
  }
  dim3 grid(ceil_div(G.n, LBN), ceil_div(G.rows, LBM), batch);
  gemm_legacy_kernel<EPI><<<grid, 256, LEG_SMEM, st>>>(G, P);
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}
template int launch_legacy<ECHO_EPI_GENERIC>(const GemmShape&, const EpiParams&, int, cudaStream_t);
template int launch_legacy<ECHO_EPI_SECTIONS>(const GemmShape&, const EpiParams&, int, cudaStream_t);
template int launch_legacy<ECHO_EPI_SWIGLU>(const GemmShape&, const EpiParams&, int, cudaStream_t);
