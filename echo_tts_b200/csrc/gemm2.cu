// bf16 GEMM / implicit causal conv, CTA-pair edition (impl 2, the product path):
//   * cluster of 2 CTAs = one 256 x BN output tile; tcgen05.mma.cta_group::2 issued by one thread of the leader
//     CTA reads A (128 rows per CTA) and B (BN/2 rows per CTA) from BOTH CTAs' shared memory, so every SM
//     stages half as many B bytes per MMA (64 B/clk/SM from L2 instead of 96) and the 192 KB ring holds 6 stages of a
//     256-wide tile (8 of a 128-wide one: operand latency, not MMA issue, bounds narrow tiles);
//   * persistent: grid = min(#pair tiles, #SM/2) clusters, static round-robin tile order with the M index
//     fastest so that the pairs working concurrently share one weight (B) tile through L2;
//   * stream-K (when the caller lends a workspace and the tile count leaves the last round ragged): the K loops of
//     all tiles form one line of k-blocks cut into equal contiguous ranges, one per pair.  A pair whose range starts
//     inside a tile dumps that raw fp32 partial accumulator to the workspace (always its FIRST unit, so it never waits
//     before publishing); the pair that owns the tile's k = 0 part (always its LAST unit) adds the partials in a
//     fixed order into its TMEM accumulator and runs the fused epilogue: deterministic, no extra kernel;
//   * two TMEM accumulator stages (2 x 256 columns): the fused epilogue of tile i runs while the MMAs of
//     tile i+1 are issued;
//   * warp roles per CTA: 0 = TMA producer (its CTA's halves; bytes land on the leader's mbarrier),
//     1 = TMEM owner (+ MMA issuer in the leader CTA), 2..9 = epilogue (two warps per TMEM lane quarter taking
//     alternate 32-column chunks; coalesced vector stores via a per-warp 32x32 shared-memory transpose and
//     auxiliary loads issued one chunk ahead, gemm_epilogue_t.cuh).
#include <stdlib.h>
#include "common.cuh"
#include "ptx_sm100.cuh"
#include "gemm_epilogue_t.cuh"

namespace {
constexpr int BM = 128;            // rows per CTA (256 per pair)
constexpr int BK = 64;
constexpr int A_BYTES = BM * BK * 2;
constexpr int B_BYTES_MAX = 128 * BK * 2;
constexpr int STAGE_BYTES = A_BYTES + B_BYTES_MAX;
constexpr int RING_BYTES = 6 * STAGE_BYTES;       // 192 KB of operand ring; narrower tiles get more, smaller stages
constexpr int MAX_STAGES = 10;
constexpr int ACC_STAGES = 2;
constexpr int ACC_COLS = 256;
constexpr int EPI_WARPS = 8;       // two per TMEM lane quarter
constexpr int B_PRODUCER_WARP = 2 + EPI_WARPS;   // optional second TMA producer (the B operand), Tiling::split_b
constexpr int NTHREADS = 64 + 32 * EPI_WARPS + 32;   // 352: register allocation rounds the CTA to 384 threads either way
constexpr int SCRATCH_BYTES = EPI_WARPS * EPI_SC_FLOATS * 4;   // per-epilogue-warp 32x32 fp32 transpose scratch
constexpr int GATE_BYTES = ACC_COLS * 4;                       // LEAN epilogue: the tile's gate row, one float per column
constexpr int SMEM2 = RING_BYTES + 1024 /*align slack*/ + 256 /*barriers*/ + SCRATCH_BYTES + GATE_BYTES;

// Debug builds only (make DEBUG=1 -> -DECHO_DEBUG): an in-kernel timeline (32 clock64 slots per CTA, enabled with
// echo_debug_gemm_trace(ptr)) and two timing experiments whose RESULTS ARE GARBAGE (operand loads left out).  None of it
// exists in the product library: the fields below are compile-time zeros there and every use folds away.
#ifdef ECHO_DEBUG
static long long* g_trace_ptr = nullptr;
static int g_debug_noload = 0;   // MMAs re-use whatever is in the ring (no TMA after the first fill): isolates MMA issue rate
static int g_debug_skip = 0;     // 1 = never load A, 2 = never load B
#else
constexpr int g_debug_noload = 0;
#endif

// n / d for 0 <= n < 2^31 with a host-made multiplier: 3 instructions instead of the ~40 of a 32-bit division (every
// thread decodes its CTA's units)
struct FastDiv {
  uint32_t mul, shift;
  __device__ __forceinline__ int div(int n) const { return (int)((__umulhi((uint32_t)n, mul) + (uint32_t)n) >> shift); }
};
static FastDiv make_fastdiv(int d) {
  FastDiv f;
  uint32_t sh = 0;
  while ((1u << sh) < (uint32_t)d) ++sh;
  f.shift = sh;
  f.mul = (uint32_t)((((uint64_t)1 << 32) * (((uint64_t)1 << sh) - (uint64_t)d)) / (uint64_t)d + 1);
  return f;
}

struct Tiling {
#ifdef ECHO_DEBUG
  long long* trace;       // nullptr unless tracing: [cta][32] SM-clock stamps
  int noload;
  int skip;               // operand whose loads are left out (1 = A, 2 = B)
#else
  static constexpr long long* trace = nullptr;
  static constexpr int noload = 0, skip = 0;
#endif
  int stages, stage_bytes; // operand ring: stage = A (16 KB) + this CTA's half of B (bn/2 rows x 128 B)
  int bn;                 // tile width (multiple of 32, <= 256); each CTA stages bn/2 rows of B
  int tiles_m, tiles_n;   // pair tiles per batch entry
  FastDiv fd_m, fd_n, fd_s;   // divisions by tiles_m / tiles_n / splits
  int num_tiles;          // work units: tiles_m * tiles_n * batch * splits
  int splits;             // split-K: units per tile (unit index = tile * splits + split)
  long long split_stride; // elements between the partial outputs
  int tail_start;         // >= 0: units from this index on are HALF-width tiles, two per wide tile (the ragged last round
                          // of a launch runs as one round of narrow tiles: shorter main loop and exposed epilogue); -1 off
  int gate_sm;            // 1: LEAN epilogue reads the gate row from shared memory (0: shuffles; A/B switch ECHO_GEMM_GATESM)
  int row_off;            // EPI_GENERIC_ROW: byte offset (from the aligned shared-memory base, inside the ring area behind the
                          // stages in use) of its staging: [residual ping-pong 8 x 2 x 4 KB when a residual is read] |
                          // bf16 chunks 8 x 2 KB | bias / alpha / 1/alpha rows 3 x 1 KB | 16 mbarriers
  int split_b;            // 1: warp B_PRODUCER_WARP issues the B-operand loads, warp 0 the A-operand loads (two issue streams)
  int stream_k;           // 1: contiguous k-block ranges per pair (num_tiles = plain tile count, splits = 1)
  int* sk_flags;          // [pairs][2 CTAs][EPI_WARPS] 0/1, self-resetting
  float* sk_partials;     // [pairs][2 CTAs][128 x 256] fp32 in the exchange layout of sk_cell()
};

constexpr int SK_FLAG_BYTES = 16384;
constexpr long long SK_SLOT_FLOATS = 2LL * BM * ACC_COLS;      // one pair's partial accumulator (256 KB)

// One work unit: k-blocks [kb0, kb1) of `tile`.  All three warp roles walk the same sequence.
// sub: -1 = full-width tile, 0 / 1 = left / right half-width tile of `tile`
// mt / nt / batch: the tile's coordinates, decoded here so that the divisions of a CTA's FIRST unit run during set-up
// (every role fetches it before the grid-dependency wait) instead of between the wait and the first operand request --
// one thread takes ~150 clocks per 32-bit and ~700 per 64-bit division, which used to put ~2.8 k clocks there.
struct Unit { int tile, split, kb0, kb1, sub, mt, nt, batch; };
struct Sched {
  int sk, total_kb, splits, unit, step, end, tail, tiles_m, tiles_n;
  FastDiv fd_m, fd_n, fd_s;
  long long u, e;
  __device__ Sched(const Tiling& T, int pair, int npairs, int total_kb_) {
    sk = T.stream_k; total_kb = total_kb_; splits = T.splits; tail = T.tail_start;
    tiles_m = T.tiles_m; tiles_n = T.tiles_n;
    fd_m = T.fd_m; fd_n = T.fd_n; fd_s = T.fd_s;
    unit = pair; step = npairs; end = T.num_tiles;
    u = e = 0;
    if (sk) {
      const long long U = (long long)T.num_tiles * total_kb;
      u = U * pair / npairs; e = U * (pair + 1) / npairs;
    }
  }
  __device__ __forceinline__ void decode(Unit& w) const {
    const int rest = fd_m.div(w.tile);
    w.mt = w.tile - rest * tiles_m;
    w.batch = fd_n.div(rest);
    w.nt = rest - w.batch * tiles_n;
  }
  __device__ bool next(Unit& w) {
    if (sk) {
      if (u >= e) return false;
      w.tile = (int)(u / total_kb); w.split = 0; w.sub = -1;
      w.kb0 = (int)(u - (long long)w.tile * total_kb);
      const long long rem = e - u;
      w.kb1 = rem < (long long)(total_kb - w.kb0) ? w.kb0 + (int)rem : total_kb;
      u += w.kb1 - w.kb0;
      decode(w);
      return true;
    }
    if (unit >= end) return false;
    if (tail >= 0 && unit >= tail) {       // narrow tail (splits == 1): two half-width units per remaining wide tile
      const int j = unit - tail;
      w.tile = tail + (j >> 1); w.sub = j & 1; w.split = 0; w.kb0 = 0; w.kb1 = total_kb;
    } else if (splits == 1) {
      w.tile = unit; w.split = 0; w.sub = -1; w.kb0 = 0; w.kb1 = total_kb;
    } else {
      w.tile = fd_s.div(unit); w.split = unit - w.tile * splits; w.sub = -1;
      w.kb0 = fd_s.div(total_kb * w.split); w.kb1 = fd_s.div(total_kb * (w.split + 1));   // 32-bit: k-blocks x splits is tiny
    }
    unit += step;
    decode(w);
    return true;
  }
};

// Exchange layout of a partial accumulator: [32-column chunk][lane quarter][4-column group][lane] float4 -- a warp
// writes / reads 512 contiguous bytes per instruction, and the reader thread is the thread position that wrote.
__device__ __forceinline__ float4* sk_cell(float* slot_cta, int chunk, int q, int c4, int lane) {
  return reinterpret_cast<float4*>(slot_cta) + (((chunk * 4 + q) * 8 + c4) * 32 + lane);
}
__device__ __forceinline__ void sk_dump(int bn, uint32_t taddr, float* slot_cta, int* flag, int q, int half, int lane) {
  for (int c = half; c < bn / 32; c += 2) {
    float v[32];
    ptx::tmem_ld32(taddr + (uint32_t)(c * 32), v);
#pragma unroll
    for (int c4 = 0; c4 < 8; ++c4)
      __stcg(sk_cell(slot_cta, c, q, c4, lane), make_float4(v[4 * c4], v[4 * c4 + 1], v[4 * c4 + 2], v[4 * c4 + 3]));
  }
  __threadfence();
  __syncwarp();
  if (lane == 0) asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(flag), "r"(1) : "memory");
}
__device__ __forceinline__ void sk_wait_flag(int* flag, int lane) {
  if (lane == 0) {
    int v = 0;
    const long long t0 = clock64();
    do {
      asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
      if (!v && clock64() - t0 > 4000000000LL) { printf("echo gemm: stream-K partial never arrived\n"); __trap(); }
    } while (!v);
  }
  __syncwarp();
  __threadfence();
}
// Sum of the partial accumulators of the pairs that continue tile `w.tile` (fixed pair order) -> the idle TMEM
// accumulator stage `tside`.  Runs on the epilogue warps while the MMAs of the owner's head unit are still in flight;
// the epilogue then reads accumulator + side stage (TmemLoader2).
__device__ __forceinline__ void sk_sideload(const Tiling& T, const Unit& w, int total_kb, int pair, int npairs, uint32_t rank,
                                            uint32_t tside, int q, int half, int lane, int ew) {
  const long long U = (long long)T.num_tiles * total_kb, tile_end = (long long)(w.tile + 1) * total_kb;
  int last = pair;
  for (int p2 = pair + 1; p2 < npairs && U * p2 / npairs < tile_end; ++p2) {
    sk_wait_flag(T.sk_flags + (p2 * 2 + (int)rank) * EPI_WARPS + ew, lane);
    last = p2;
  }
  float* const slot0 = T.sk_partials + (long long)rank * (BM * ACC_COLS);
  const int nch = T.bn / 32;
  // two chunks (64 registers) of every contributor are requested before anything is consumed: two memory round trips
  // per warp instead of one per chunk (the asm statements around TMEM accesses are compiler barriers)
#pragma unroll 1
  for (int i0 = 0; i0 < 4; i0 += 2) {
    float4 d[2][8];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int c = half + 2 * (i0 + i);
      if (c < nch) {
#pragma unroll
        for (int c4 = 0; c4 < 8; ++c4) d[i][c4] = __ldcg(sk_cell(slot0 + (long long)(pair + 1) * SK_SLOT_FLOATS, c, q, c4, lane));
      }
    }
    for (int p2 = pair + 2; p2 <= last; ++p2) {
      float* slot = slot0 + (long long)p2 * SK_SLOT_FLOATS;
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int c = half + 2 * (i0 + i);
        if (c < nch) {
          float4 t[8];
#pragma unroll
          for (int c4 = 0; c4 < 8; ++c4) t[c4] = __ldcg(sk_cell(slot, c, q, c4, lane));
#pragma unroll
          for (int c4 = 0; c4 < 8; ++c4) {
            d[i][c4].x += t[c4].x; d[i][c4].y += t[c4].y; d[i][c4].z += t[c4].z; d[i][c4].w += t[c4].w;
          }
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int c = half + 2 * (i0 + i);
      if (c < nch) ptx::tmem_st32(tside + (uint32_t)(c * 32), reinterpret_cast<const float*>(d[i]));
    }
  }
  ptx::tmem_st_wait();
  __syncwarp();
  if (lane == 0)
    for (int p2 = pair + 1; p2 <= last; ++p2)        // consumed: ready for the next launch
      *reinterpret_cast<volatile int*>(T.sk_flags + (p2 * 2 + (int)rank) * EPI_WARPS + ew) = 0;
}

struct TmemLoader2 {
  uint32_t taddr;
  __device__ __forceinline__ void load32(int col, float* v) { ptx::tmem_ld32(taddr + (uint32_t)col, v); }
  __device__ __forceinline__ void issue32(int col, float* v) { ptx::tmem_ld32_issue(taddr + (uint32_t)col, v); }
  __device__ __forceinline__ void wait() { ptx::tmem_ld_wait(); }
};
// accumulator += side stage (TMEM -> registers -> TMEM), this warp's chunks
__device__ __forceinline__ void sk_merge(int bn, uint32_t tacc, uint32_t tside, int half) {
  for (int c = half; c < bn / 32; c += 2) {
    float v[32], s[32];
    ptx::tmem_ld32(tacc + (uint32_t)(c * 32), v);
    ptx::tmem_ld32(tside + (uint32_t)(c * 32), s);
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] += s[i];
    ptx::tmem_st32(tacc + (uint32_t)(c * 32), v);
  }
  ptx::tmem_st_wait();
}

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NTHREADS, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_bn, const __grid_constant__ CUtensorMap tmap_out,
                 const __grid_constant__ CUtensorMap tmap_res, const __grid_constant__ CUtensorMap tmap_ob,
                 const __grid_constant__ GemmShape G, const __grid_constant__ EpiParams P, const __grid_constant__ Tiling T) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  // operand ring | per-warp epilogue scratch (1024-byte aligned: the TMA epilogue reads it as SWIZZLE_128B tiles) | barriers
  float* scratch = reinterpret_cast<float*>(smem + RING_BYTES);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + RING_BYTES + SCRATCH_BYTES);
  uint64_t* empty_bar = full_bar + MAX_STAGES;
  uint64_t* tmem_full_bar = empty_bar + MAX_STAGES;
  uint64_t* tmem_empty_bar = tmem_full_bar + ACC_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + ACC_STAGES);
  float* gate_sm = reinterpret_cast<float*>(smem + RING_BYTES + SCRATCH_BYTES + 256);
  const int STAGES = T.stages, STAGE_SZ = T.stage_bytes;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
  const int total_kb = G.k_blocks_per_tap * G.taps;
  const int half_n = T.bn >> 1;
#ifdef ECHO_DEBUG
  long long* const trace = T.trace ? T.trace + (long long)blockIdx.x * 32 : nullptr;
#else
  constexpr long long* trace = nullptr;
#endif
  if (trace && threadIdx.x == 0) trace[0] = clock64();
  // every role walks the same unit sequence; the first unit (and its divisions) is ready before the dependency wait
  Sched sch(T, pair, npairs, total_kb);
  Unit w;
  bool have = sch.next(w);

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_a);
    ptx::prefetch_tensormap(&tmap_b);
    if (T.tail_start >= 0) ptx::prefetch_tensormap(&tmap_bn);
    if (EPI == EPI_GENERIC_LEAN || EPI == EPI_GENERIC_ROW) ptx::prefetch_tensormap(&tmap_out);
    if (EPI == EPI_GENERIC_ROW) { ptx::prefetch_tensormap(&tmap_res); ptx::prefetch_tensormap(&tmap_ob); }
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(&full_bar[s], 1);        // leader's producer: arrive.expect_tx (both CTAs' bytes)
      ptx::mbar_init(&empty_bar[s], 1);       // tcgen05.commit multicast to both CTAs
    }
    for (int a = 0; a < ACC_STAGES; ++a) {
      ptx::mbar_init(&tmem_full_bar[a], 1);   // tcgen05.commit multicast to both CTAs
      ptx::mbar_init(&tmem_empty_bar[a], 2 * EPI_WARPS);  // every epilogue warp of both CTAs arrives on the leader's barrier
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc_cg2(tmem_slot, ACC_STAGES * ACC_COLS);
    ptx::tmem_relinquish_cg2();
  }
  ptx::tc_fence_before();
  ptx::cluster_sync_all();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (trace && threadIdx.x == 0) trace[1] = clock64();
  pdl_launch_dependents();    // the next kernel may begin its own setup
  // The grid dependency is waited for PER ROLE, as late as each role allows: the A producer immediately in front of its
  // first load (scheduling arithmetic, barrier bookkeeping and the instruction fetches of the loop head are done by then:
  // they used to cost ~1.2 k clocks between the wait and the first request), the B producer never when B is a load-time
  // tensor (its first tiles stream from HBM while the previous launch drains), the MMA thread never (no global access),
  // the epilogue warps before their first global access.

  // TMA producer loop.  do_a / do_b: which operand this thread loads; the leader CTA's A producer also posts the
  // transaction count of the whole stage (both CTAs, both operands) on the leader's full barrier.
  auto produce = [&](const bool do_a, const bool do_b) {
    int s = 0;
    uint32_t ph = 0;
    bool first_unit = true;
    bool must_wait = do_a || !G.b_static;
    for (; have; have = sch.next(w)) {
      const int batch = w.batch;
      const int m0 = w.mt * (2 * BM) + (int)rank * BM;
      const int half_u = w.sub < 0 ? half_n : half_n >> 1;            // rows of B this CTA stages for the unit
      const int n0 = w.nt * T.bn + (w.sub > 0 ? half_n : 0) + (int)rank * half_u;
      const uint32_t stage_tx = 2u * (uint32_t)(A_BYTES + half_u * BK * 2);
      const int kb0 = w.kb0, kb1 = w.kb1;
      const bool first_u = first_unit;
      first_unit = false;
      int tap = 0, kk = kb0;                                          // k-block kb = (tap, kk), tracked without dividing
      if (G.taps > 1) { tap = kb0 / G.k_blocks_per_tap; kk = kb0 - tap * G.k_blocks_per_tap; }
      for (int kb = kb0; kb < kb1; ++kb, ++kk) {
        if (T.noload && (kb - kb0 >= STAGES || !first_u)) break;
        if (kk == G.k_blocks_per_tap) { kk = 0; ++tap; }
        ptx::mbar_wait(&empty_bar[s], ph ^ 1);
        uint8_t* sa = smem + s * STAGE_SZ;
        uint8_t* sb = sa + A_BYTES;
        const uint32_t tx = T.skip == 1 ? 2u * (uint32_t)(half_u * BK * 2) : T.skip == 2 ? 2u * (uint32_t)A_BYTES : stage_tx;
        if (do_a && rank == 0) ptx::mbar_arrive_expect_tx(&full_bar[s], tx);
        const uint32_t lbar = ptx::peer0_addr(&full_bar[s]);
        if (must_wait) {
          must_wait = false;
          pdl_wait();
          if (trace && do_a && !T.stream_k) trace[3] = clock64();    // grid dependency resolved (A producer's view)
        }
        if (do_a && T.skip != 1) ptx::tma_load_3d_cg2(sa, &tmap_a, lbar, kk * BK, m0 + G.tap_shift[tap], batch);
        if (do_b && T.skip != 2) ptx::tma_load_2d_cg2(sb, w.sub < 0 ? &tmap_b : &tmap_bn, lbar, tap * G.k_tap + kk * BK, n0);
        if (trace && do_a && first_u && kb == kb0 && !T.stream_k) trace[4] = clock64();   // first stage requested
        if (++s == STAGES) { s = 0; ph ^= 1; }
      }
    }
  };
  if (warp == 0) {
    if (lane == 0) produce(true, T.split_b == 0);
    __syncwarp();
  } else if (warp == B_PRODUCER_WARP) {
    // B bytes may land on the leader's full barrier before the A producer has posted the stage's transaction count: the
    // phase cannot complete until that arrive, and a transiently negative tx-count is legal
    if (T.split_b && lane == 0) produce(false, true);
    __syncwarp();
  } else if (warp == 1) {
    if (rank == 0) {   // whole warp converged; one elected lane issues (keeps UTCHMMA free of election loops)
      const uint32_t idesc_wide = ptx::umma_idesc_bf16(2 * BM, T.bn), idesc_narrow = ptx::umma_idesc_bf16(2 * BM, T.bn >> 1);
      uint32_t tc = 0, ph = 0;
      int s = 0;
      bool first = true, first_unit = true;
      for (; have; have = sch.next(w), ++tc) {
        const int kb0 = w.kb0, kb1 = w.kb1;
        const uint32_t idesc = w.sub < 0 ? idesc_wide : idesc_narrow;
        const bool first_u = first_unit;
        first_unit = false;
        const uint32_t as = tc & 1, aph = (tc >> 1) & 1;
        if (lane == 0) ptx::mbar_wait(&tmem_empty_bar[as], aph ^ 1);     // epilogues of both CTAs drained this accumulator
        __syncwarp();
        ptx::tc_fence_after();
        if (trace && tc < 6 && lane == 0) trace[8 + tc * 4 + 0] = clock64();
        const uint32_t tacc = tmem_base + as * ACC_COLS;
        for (int kb = kb0; kb < kb1; ++kb) {
          const uint32_t sa = smem_u32(smem + s * STAGE_SZ);
          const uint64_t da = ptx::umma_desc_sw128(sa);
          const uint64_t db = ptx::umma_desc_sw128(sa + A_BYTES);
          if (ptx::elect_one()) {      // the elected lane alone polls the barrier (31 spinning lanes only steal issue slots)
            if (!(T.noload && (kb - kb0 >= STAGES || !first_u))) ptx::mbar_wait(&full_bar[s], ph);
            ptx::tc_fence_after();
            if (trace && first) trace[2] = clock64();
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              ptx::umma_bf16_cg2(tacc, da + 2 * k, db + 2 * k, idesc, (kb != kb0 || k != 0) ? 1u : 0u);
            if (!T.noload) ptx::umma_commit_cg2_mc(&empty_bar[s], 0b11);   // both CTAs may refill this slot
          }
          __syncwarp();
          first = false;
          if (++s == STAGES) { s = 0; ph ^= 1; }
        }
        if (ptx::elect_one()) ptx::umma_commit_cg2_mc(&tmem_full_bar[as], 0b11);
        __syncwarp();
        if (trace && tc < 6 && lane == 0) trace[8 + tc * 4 + 1] = clock64();
      }
    }
    __syncwarp();
  } else {
    pdl_wait();                              // bias / gate / residual reads and every store below
    const int q = warp & 3;                  // TMEM lane quarter this warp may read
    const int ew = warp - 2, half = ew >> 2;
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    uint32_t tc = 0;
    auto release_acc = [&](uint32_t as) {
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(ptx::peer0_addr(&tmem_empty_bar[as]));
    };
    RowStage RS;
    if constexpr (EPI == EPI_GENERIC_ROW) {
      uint8_t* base = smem + T.row_off;
      const bool has_res = P.resid != nullptr;
      uint8_t* after_res = base + (has_res ? EPI_WARPS * 2 * 4096 : 0);
      RS.out = scratch + ew * EPI_SC_FLOATS;
      RS.res[0] = reinterpret_cast<float*>(base + ew * 8192);
      RS.res[1] = RS.res[0] + 1024;
      RS.outb = reinterpret_cast<uint32_t*>(after_res + ew * 2048);
      RS.rows = reinterpret_cast<const float*>(after_res + EPI_WARPS * 2048);
      RS.res_bar = reinterpret_cast<uint64_t*>(after_res + EPI_WARPS * 2048 + 3 * ACC_COLS * 4) + 2 * ew;
      RS.res_phase = 0;
      RS.ld_rows = ACC_COLS;
      if (lane == 0) {
        ptx::mbar_init(&RS.res_bar[0], 1);
        ptx::mbar_init(&RS.res_bar[1], 1);
        ptx::fence_barrier_init();
      }
      __syncwarp();
    }
    // The stream-K special cases are peeled off the steady-state loop (first unit = dump, last unit = head) so that
    // their register needs do not lengthen live ranges inside it.
    if (have && T.stream_k && w.kb0 > 0) {
      // the tile's k = 0 part belongs to an earlier pair: publish this partial accumulator, no epilogue
      ptx::mbar_wait(&tmem_full_bar[0], 0);
      ptx::tc_fence_after();
      if (trace && warp == 2 && lane == 0) trace[10] = clock64();
      sk_dump(T.bn, tmem_base + lane_base, T.sk_partials + (long long)pair * SK_SLOT_FLOATS + (long long)rank * (BM * ACC_COLS),
              T.sk_flags + (pair * 2 + (int)rank) * EPI_WARPS + ew, q, half, lane);
      release_acc(0);
      if (trace && warp == 2 && lane == 0) trace[11] = clock64();
      ++tc;
      have = sch.next(w);
    }
    for (; have && !(T.stream_k && w.kb1 < total_kb); have = sch.next(w), ++tc) {
      const int split = w.split;
      const int nt = w.nt, batch = w.batch;
      const int m0 = w.mt * (2 * BM) + (int)rank * BM;
      const uint32_t as = tc & 1, aph = (tc >> 1) & 1;
      const int bn_u = w.sub < 0 ? T.bn : T.bn >> 1;
      const int nc0 = nt * T.bn + (w.sub > 0 ? (T.bn >> 1) : 0);
      EpiPre<EPI> pre;
      epilogue_prefetch_t<EPI>(bn_u, P, lane, half, batch, m0 + q * 32, nc0, pre);
      // LEAN: when the CTA's 128 rows share ONE gate row (CTA-uniform test), the eight warps stage that row's bn_u values
      // in shared memory -- a broadcast LDS.128 per four columns replaces 32 shuffles per chunk in the exposed epilogue.
      // Barrier in front: the previous tile's readers are done; barrier behind: the row is complete.
      const float* gsm = nullptr;
      if (EPI == EPI_GENERIC_LEAN && T.gate_sm && P.gate && m0 < P.rows) {
        const unsigned rpg = (unsigned)P.rows_per_gate;
        const unsigned mg_first = (unsigned)batch * (unsigned)P.rows + (unsigned)m0;
        const unsigned mg_last = mg_first + (unsigned)(min(BM, P.rows - m0) - 1);
        const unsigned g_first = mg_first / rpg;
        if (g_first == mg_last / rpg) {
          asm volatile("bar.sync 3, %0;" ::"n"(32 * EPI_WARPS) : "memory");
          const int col = ew * 32 + lane;
          if (col < bn_u) gate_sm[col] = nc0 + col < P.n ? __ldg(P.gate + (long long)g_first * P.gate_ld + nc0 + col) : 0.f;
          asm volatile("bar.sync 3, %0;" ::"n"(32 * EPI_WARPS) : "memory");
          gsm = gate_sm;
        }
      }
      if constexpr (EPI == EPI_GENERIC_ROW) {
        // the tile's bias / Snake vectors as shared-memory rows (broadcast reads in the row-per-thread layout).  Barrier in
        // front: the previous tile's readers are done; barrier behind: the rows are complete.
        float* rows = const_cast<float*>(RS.rows);
        asm volatile("bar.sync 3, %0;" ::"n"(32 * EPI_WARPS) : "memory");
        const int col = ew * 32 + lane, n = nc0 + col;
        const bool ok = col < bn_u && n < P.n;
        rows[col] = (ok && P.bias) ? __ldg(P.bias + n) : 0.f;
        if (P.out2_bf16) {
          const float al = ok ? __ldg(P.snake_alpha + n) : 1.f;
          rows[ACC_COLS + col] = al;
          rows[2 * ACC_COLS + col] = __fdividef(1.f, al + 1e-9f);
        }
        asm volatile("bar.sync 3, %0;" ::"n"(32 * EPI_WARPS) : "memory");
      }
      ptx::mbar_wait(&tmem_full_bar[as], aph);
      ptx::tc_fence_after();
      if (trace && tc < 6 && warp == 2 && lane == 0) trace[8 + tc * 4 + 2] = clock64();
      TmemLoader2 L{tmem_base + as * ACC_COLS + lane_base};
      // LEAN, last unit of this CTA: every MMA of the launch that reads this CTA's ring has retired (the accumulator is
      // complete and no load follows), so the ring stages each chunk in its own 4 KB slot -- no waiting between chunks
      float* stage = scratch + ew * EPI_SC_FLOATS;
      int stage_stride = 0;
      if (EPI == EPI_GENERIC_LEAN && !T.stream_k) {
        Sched peek = sch;
        Unit wn;
        if (!peek.next(wn)) {
          stage = reinterpret_cast<float*>(smem) + ew * (ACC_COLS / 64) * EPI_SC_FLOATS;
          stage_stride = EPI_SC_FLOATS;
        }
      }
      if constexpr (EPI == EPI_GENERIC_ROW) {
        const EpiRows R = epi_rows(P, half, batch, m0 + q * 32, nc0, 0, 0);
        epi_t_row(bn_u, P, L, lane, R, RS, &tmap_out, &tmap_res, &tmap_ob);
      } else {
        run_epilogue_t<EPI>(bn_u, P, L, stage, lane, half, batch, m0 + q * 32, nc0, pre, &tmap_out, split,
                            split * T.split_stride, stage_stride, gsm);
      }
      release_acc(as);
      if (trace && tc < 6 && warp == 2 && lane == 0) trace[8 + tc * 4 + 3] = clock64();
    }
    if (have) {
      // stream-K head: k-blocks [0, kb1) of a tile that later pairs continue (always this pair's last unit).  The other
      // accumulator stage is idle from here on: fill it with their partial sum while our MMAs still run, merge, then
      // run the fused epilogue on the complete accumulator.
      const int nt = w.nt, batch = w.batch;
      const int m0 = w.mt * (2 * BM) + (int)rank * BM;
      const uint32_t as = tc & 1, aph = (tc >> 1) & 1;
      const uint32_t tacc = tmem_base + as * ACC_COLS + lane_base, tside = tmem_base + (as ^ 1u) * ACC_COLS + lane_base;
      if (trace && warp == 2 && lane == 0) trace[3] = clock64();
      sk_sideload(T, w, total_kb, pair, npairs, rank, tside, q, half, lane, ew);
      if (trace && warp == 2 && lane == 0) trace[4] = clock64();
      ptx::mbar_wait(&tmem_full_bar[as], aph);
      ptx::tc_fence_after();
      if (trace && tc < 6 && warp == 2 && lane == 0) trace[8 + tc * 4 + 2] = clock64();
      sk_merge(T.bn, tacc, tside, half);
      ptx::tc_fence_before();
      asm volatile("bar.sync 2, %0;" ::"n"(32 * EPI_WARPS) : "memory");   // fused epilogues read neighbouring chunks
      ptx::tc_fence_after();
      if (trace && warp == 2 && lane == 0) trace[5] = clock64();
      TmemLoader2 L{tacc};
      EpiPre<EPI> pre;
      epilogue_prefetch_t<EPI>(T.bn, P, lane, half, batch, m0 + q * 32, nt * T.bn, pre);
      if constexpr (EPI != EPI_GENERIC_ROW)     // (the row epilogue is never scheduled stream-K)
        run_epilogue_t<EPI>(T.bn, P, L, scratch + ew * EPI_SC_FLOATS, lane, half, batch, m0 + q * 32, nt * T.bn, pre,
                            &tmap_out, 0, 0);
      release_acc(as);
      if (trace && tc < 6 && warp == 2 && lane == 0) trace[8 + tc * 4 + 3] = clock64();
    }
  }
  ptx::tc_fence_before();
  if (trace && threadIdx.x == 0) trace[6] = clock64();
  ptx::cluster_sync_all();
  if (warp == 1) ptx::tmem_dealloc_cg2(tmem_base, ACC_STAGES * ACC_COLS);
  if (trace && threadIdx.x == 0) trace[7] = clock64();
}

int getenv_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e && e[0] ? atoi(e) : dflt;
}
// scheduler switches (A/B measurements), read ONCE per process -- not per launch
int use_narrow_tail() { static const int v = getenv_int("ECHO_GEMM_TAIL", 1); return v; }
int use_gate_sm() { static const int v = getenv_int("ECHO_GEMM_GATESM", 1); return v; }
int use_row_epilogue() { static const int v = getenv_int("ECHO_GEMM_ROWEPI", 1); return v; }
bool row_epilogue_aligned(const EpiParams& P) {
  auto al16 = [](const void* p) { return ((uintptr_t)p % 16) == 0; };
  return al16(P.out_f32) && al16(P.resid) && al16(P.out_bf16) && al16(P.out2_bf16) && P.ld_out_f32 % 4 == 0 &&
         P.ld_resid % 4 == 0 && P.ld_out_bf16 % 8 == 0 && P.ld_out2_bf16 % 8 == 0;
}
int use_split_b() { static const int v = getenv_int("ECHO_GEMM_SPLITB", 1); return v; }
int max_stages() { static const int v = getenv_int("ECHO_GEMM_STAGES", MAX_STAGES); return v < 2 ? 2 : v; }
int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}
}  // namespace

int echo_make_tmap_bf16(CUtensorMap* m, const void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides_b,
                        const cuuint32_t* box);   // gemm.cu
int echo_make_tmap_f32(CUtensorMap* m, const void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides_b,
                       const cuuint32_t* box);    // gemm.cu
int echo_make_tmap_bf16_linear(CUtensorMap* m, const void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides_b,
                               const cuuint32_t* box);   // gemm.cu

// Bytes of the optional stream-K workspace (echo_gemm_desc.workspace): flags + one partial accumulator per CTA pair.
// The caller zeroes it once after allocation and must not share it between streams that run GEMMs concurrently.
extern "C" long long echo_gemm_workspace_bytes(void) {
  return SK_FLAG_BYTES + (long long)(num_sms() / 2) * SK_SLOT_FLOATS * (long long)sizeof(float);
}

template <int EPI>
static int launch_pair_t(int bn, int splits, long long split_stride, void* workspace, long long workspace_bytes,
                         const GemmShape& G, const EpiParams& P, int batch, cudaStream_t st, int sched_flags) {
  CUtensorMap ta, tb, tbn;
  // Narrow tail: when the last round of whole-tile scheduling is ragged and its tiles, cut in two along N, still fit ONE
  // round, those tiles run half-width -- a shorter main loop and half the exposed epilogue at the end of the launch
  // (q/k/v/gate projection: 256 tiles = 3 rounds of 74 + 34 -> 68 half tiles; 96 tiles at 640 tokens = 74 + 22 -> 44).
  int tail_start = -1;
  {
    const int P2 = num_sms() / 2;
    const int wide = ceil_div(G.rows, 2 * BM) * ceil_div(G.n, bn) * batch;
    const int rem = wide % P2;
    if (splits == 1 && bn == 256 && G.n % 256 == 0 && !g_debug_noload && use_narrow_tail() && !(sched_flags & 2)) {
      if (wide > P2 && rem > 0 && 2 * rem <= P2) tail_start = wide - rem;
      else if (2 * wide <= P2) tail_start = 0;      // a single sparse round: every tile half-width, twice the pairs busy
    }
  }
  {
    cuuint64_t dims[3] = {(cuuint64_t)G.k_tap, (cuuint64_t)(G.rows + G.a_ctx), (cuuint64_t)batch};
    cuuint64_t str[2] = {(cuuint64_t)G.lda * 2, (cuuint64_t)(batch > 1 ? G.a_batch_stride : (long long)(G.rows + G.a_ctx) * G.lda) * 2};
    cuuint32_t box[3] = {BK, BM, 1};
    int rc = echo_make_tmap_bf16(&ta, G.a, 3, dims, str, box);
    if (rc) return rc;
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)G.taps * G.k_tap, (cuuint64_t)G.n};
    cuuint64_t str[1] = {(cuuint64_t)G.ldb * 2};
    cuuint32_t box[2] = {BK, (cuuint32_t)(bn / 2)};
    int rc = echo_make_tmap_bf16(&tb, G.b, 2, dims, str, box);
    if (rc) return rc;
    tbn = tb;                                   // half-width tiles of the narrow tail: a box of bn/4 rows
    if (tail_start >= 0) {
      box[1] = (cuuint32_t)(bn / 4);
      rc = echo_make_tmap_bf16(&tbn, G.b, 2, dims, str, box);
      if (rc) return rc;
    }
  }
  CUtensorMap tout = ta;                        // unused unless the epilogue leaves through TMA
  if (EPI == EPI_GENERIC_LEAN) {
    // fp32 output [splits][batch][rows][n]: 32 x 32 boxes; rows past a batch entry / columns past n are clipped
    const long long ld = P.ld_out_f32;
    const long long bstride = (long long)G.rows * ld;
    const long long sstride = splits > 1 ? split_stride : bstride * batch;
    cuuint64_t dims[4] = {(cuuint64_t)G.n, (cuuint64_t)G.rows, (cuuint64_t)batch, (cuuint64_t)splits};
    cuuint64_t str[3] = {(cuuint64_t)ld * 4, (cuuint64_t)bstride * 4, (cuuint64_t)sstride * 4};
    cuuint32_t box[4] = {32, 32, 1, 1};
    int rc = echo_make_tmap_f32(&tout, P.out_f32, 4, dims, str, box);
    if (rc) return rc;
  }
  CUtensorMap tres = ta, tob = ta;              // row epilogue: residual (fp32, loaded) and bf16 output (stored) tiles
  if (EPI == EPI_GENERIC_ROW) {
    cuuint32_t box[4] = {32, 32, 1, 1};
    if (P.out_f32) {
      const long long ld = P.ld_out_f32;
      cuuint64_t dims[4] = {(cuuint64_t)G.n, (cuuint64_t)G.rows, (cuuint64_t)batch, 1};
      cuuint64_t str[3] = {(cuuint64_t)ld * 4, (cuuint64_t)G.rows * ld * 4, (cuuint64_t)G.rows * ld * batch * 4};
      if (int rc = echo_make_tmap_f32(&tout, P.out_f32, 4, dims, str, box)) return rc;
    }
    if (P.resid) {
      const long long ld = P.ld_resid;
      cuuint64_t dims[4] = {(cuuint64_t)G.n, (cuuint64_t)G.rows, (cuuint64_t)batch, 1};
      cuuint64_t str[3] = {(cuuint64_t)ld * 4, (cuuint64_t)G.rows * ld * 4, (cuuint64_t)G.rows * ld * batch * 4};
      if (int rc = echo_make_tmap_f32(&tres, P.resid, 4, dims, str, box)) return rc;
    }
    if (P.out2_bf16 || P.out_bf16) {
      const bf16* ob = P.out2_bf16 ? P.out2_bf16 : P.out_bf16;
      const long long ld = P.out2_bf16 ? P.ld_out2_bf16 : P.ld_out_bf16;
      cuuint64_t dims[4] = {(cuuint64_t)G.n, (cuuint64_t)G.rows, (cuuint64_t)batch, 1};
      cuuint64_t str[3] = {(cuuint64_t)ld * 2, (cuuint64_t)G.rows * ld * 2, (cuuint64_t)G.rows * ld * batch * 2};
      if (int rc = echo_make_tmap_bf16_linear(&tob, ob, 4, dims, str, box)) return rc;
    }
  }
  {   // the attribute is per device: a process that drives several GPUs sets it once on each
    static bool attr_set[64] = {};
    int dev = 0;
    ECHO_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
      ECHO_CUDA(cudaFuncSetAttribute(gemm_pair_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM2));
      if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
  }
  Tiling T;
  T.split_b = use_split_b();
  T.gate_sm = use_gate_sm();
#ifdef ECHO_DEBUG
  T.trace = g_trace_ptr;
  T.noload = g_debug_noload;
  T.skip = g_debug_skip;
#endif
  T.bn = bn;
  T.stage_bytes = A_BYTES + (bn / 2) * BK * 2;
  T.stages = RING_BYTES / T.stage_bytes < MAX_STAGES ? RING_BYTES / T.stage_bytes : MAX_STAGES;
  if (T.stages > max_stages()) T.stages = max_stages();
  T.row_off = 0;
  if (EPI == EPI_GENERIC_ROW) {                 // its staging lives in the ring area, behind the stages in use
    const int extra = (P.resid ? EPI_WARPS * 2 * 4096 : 0) + EPI_WARPS * 2048 + 3 * ACC_COLS * 4 + 2 * EPI_WARPS * 8;
    const int fit = (RING_BYTES - extra) / T.stage_bytes;
    if (T.stages > fit) T.stages = fit;
    T.row_off = (T.stages * T.stage_bytes + 1023) & ~1023;
    if (T.stages < 2 || T.row_off + extra > RING_BYTES) {
      echo_set_error("pair gemm: row epilogue staging does not fit (block_n %d)", bn);
      return ECHO_ERR_UNSUPPORTED;
    }
  }
  T.tiles_m = ceil_div(G.rows, 2 * BM);
  T.tiles_n = ceil_div(G.n, bn);
  T.fd_m = make_fastdiv(T.tiles_m); T.fd_n = make_fastdiv(T.tiles_n); T.fd_s = make_fastdiv(splits);
  T.splits = splits;
  T.split_stride = split_stride;
  T.num_tiles = T.tiles_m * T.tiles_n * batch * splits;
  T.tail_start = tail_start;
  if (tail_start >= 0) T.num_tiles = tail_start + 2 * (T.num_tiles - tail_start);
  int pairs = T.num_tiles < num_sms() / 2 ? T.num_tiles : num_sms() / 2;
  T.stream_k = 0; T.sk_flags = nullptr; T.sk_partials = nullptr;
  if (splits == 1 && tail_start < 0 && workspace && workspace_bytes >= echo_gemm_workspace_bytes() && !g_debug_noload) {
    // stream-K when whole-tile scheduling would idle more than ~8 % of the machine in its last round and every pair
    // still gets a few k-blocks of work
    const int P2 = num_sms() / 2, total_kb = G.k_blocks_per_tap * G.taps;
    const long long tiles = T.num_tiles, rounds = (tiles + P2 - 1) / P2;
    const bool ragged = tiles * 100 < rounds * P2 * 92;
    if (ragged && total_kb >= 8 && tiles * total_kb >= 4LL * P2 && (sched_flags & 1)) {
      T.stream_k = 1;
      T.sk_flags = reinterpret_cast<int*>(workspace);
      T.sk_partials = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + SK_FLAG_BYTES);
      pairs = P2;
    }
  }
  ECHO_CUDA(echo_launch(gemm_pair_kernel<EPI>, dim3(2 * pairs), dim3(NTHREADS), SMEM2, st, ta, tb, tbn, tout, tres, tob, G, P, T));
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}

// Debug aid: device buffer of >= 32 * grid int64 that the next pair-GEMM launches fill with clock64 stamps:
// [0] start, [1] setup done, [2] first operands landed, [6] all roles done, [7] exit, and for the CTA's first six
// tiles t: [8+4t] accumulator stage free (MMA thread), [9+4t] MMAs issued, [10+4t] accumulator ready (epilogue
// warp 2), [11+4t] epilogue done.  Pass NULL to switch tracing off.
extern "C" int echo_debug_gemm_trace(long long* device_buf) {
#ifdef ECHO_DEBUG
  g_trace_ptr = device_buf;
  const char* e = getenv("ECHO_GEMM_NOLOAD");    // results are garbage in this mode; timing experiments only
  g_debug_noload = (device_buf && e && e[0] == '1') ? 1 : 0;
  const char* k = getenv("ECHO_GEMM_SKIP");      // "A" / "B": leave that operand's loads out (timing experiments only)
  g_debug_skip = (device_buf && k) ? (k[0] == 'A' ? 1 : k[0] == 'B' ? 2 : 0) : 0;
  // lean-epilogue experiments (bitmask, see gemm_epilogue_t.cuh) and its per-chunk stamps: the buffer's second half
  // ([grid .. 2 grid) x 32 slots) when ECHO_GEMM_EPITRACE=<grid> is set
  const int dbg = device_buf ? getenv_int("ECHO_GEMM_EPIDBG", 0) : 0;
  const int egrid = device_buf ? getenv_int("ECHO_GEMM_EPITRACE", 0) : 0;
  long long* etr = egrid > 0 ? device_buf + (long long)egrid * 32 : nullptr;
  ECHO_CUDA(cudaMemcpyToSymbol(g_epi_dbg, &dbg, sizeof(int)));
  ECHO_CUDA(cudaMemcpyToSymbol(g_epi_trace, &etr, sizeof(etr)));
  return ECHO_OK;
#else
  if (!device_buf) return ECHO_OK;
  echo_set_error("echo_debug_gemm_trace: this library was built without -DECHO_DEBUG (make -C echo_tts_b200/csrc DEBUG=1)");
  return ECHO_ERR_UNSUPPORTED;
#endif
}

// bn: multiple of 32 in [64, 256] (so that each CTA's half is a legal UMMA_N/2 and a whole number of 16-column groups)
int echo_launch_gemm_pair(int epi, int bn, int splits, long long split_stride, void* workspace, long long workspace_bytes,
                          const GemmShape& G, const EpiParams& P, int batch, cudaStream_t st, int sched_flags) {
  if (splits < 1) splits = 1;
  if (splits > 1 && (epi != ECHO_EPI_GENERIC || splits > G.k_blocks_per_tap * G.taps)) {
    echo_set_error("pair gemm: k_splits %d needs the GENERIC epilogue and at least one k-block per split", splits);
    return ECHO_ERR_UNSUPPORTED;
  }
  if (bn < 64 || bn > 256 || bn % 32 != 0) {
    echo_set_error("pair gemm: block_n %d must be a multiple of 32 in [64, 256]", bn);
    return ECHO_ERR_UNSUPPORTED;
  }
  switch (epi) {
    case ECHO_EPI_GENERIC:
      // LEAN (TMA epilogue): fp32 output only, residual absent or updated in place
      if (P.act == ECHO_ACT_NONE && !P.out_bf16 && !P.out2_bf16 && P.out_f32 && (P.red_resid || !P.resid))
        return launch_pair_t<EPI_GENERIC_LEAN>(bn, splits, split_stride, workspace, workspace_bytes, G, P, batch, st, sched_flags);
      // ROW (TMA in and out, row per thread): bias [+ residual that is read] -> fp32 and / or one bf16 output (plain or
      // Snake) -- the S1-DAC convolutions; needs TMA-legal bases and pitches
      if (use_row_epilogue() && splits == 1 && P.act == ECHO_ACT_NONE && !P.gate && !P.red_resid &&
          !(P.out_bf16 && P.out2_bf16) && (P.out_f32 || P.out_bf16 || P.out2_bf16) && !(sched_flags & 5) &&
          row_epilogue_aligned(P))
        return launch_pair_t<EPI_GENERIC_ROW>(bn, 1, 0, nullptr, 0, G, P, batch, st, sched_flags);
      return launch_pair_t<ECHO_EPI_GENERIC>(bn, splits, split_stride, workspace, workspace_bytes, G, P, batch, st, sched_flags);
    case ECHO_EPI_SECTIONS: return launch_pair_t<ECHO_EPI_SECTIONS>(bn, 1, 0, workspace, workspace_bytes, G, P, batch, st, sched_flags);
    default: return launch_pair_t<ECHO_EPI_SWIGLU>(bn, 1, 0, workspace, workspace_bytes, G, P, batch, st, sched_flags);
  }
}
