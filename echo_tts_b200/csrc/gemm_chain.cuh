// Two dependent GEMMs in ONE persistent launch of the CTA-pair kernel (included by gemm2.cu): the DiT's MLP,
//   phase 0:  h = silu(xn w1^T) * (xn w3^T)            (SwiGLU epilogue, bf16 h; model.py:308)
//   phase 1:  x += gate * (h w2^T)  or its split-K partials   (lean TMA epilogue; model.py:388)
// Run as two launches, each GEMM pays its own set-up, pipeline fill, ragged last round and exposed last epilogue
// (measured: 7.4 + 3.8 us of 97 at 1920 tokens, 4.8 + 6.8 us of 44 at 640).  Here the pairs walk ONE unit list -- all
// phase-0 tiles, then all phase-1 units -- and the operand ring, the two TMEM accumulator stages and the epilogue warps
// never drain in between: the last SwiGLU epilogues run under the first MMAs of the down-projection.
//
// The dependency is per 256-row block, not grid-wide: phase-0 tiles are ordered row block first (N index fastest), every
// epilogue warp counts itself into ready[row block] after its stores (release), and the A producer of a phase-1 unit
// polls that counter (acquire, then a generic -> async proxy fence in front of the TMA reads of h) before its first load.
// B (weights) never waits, so the ring is already full of w2 when h becomes readable.  With static round-robin
// assignment a pair reaches its first phase-1 unit after every phase-0 tile of that row block has been ISSUED by some
// pair (they precede it in the list and no phase-0 tile ever waits), so the wait is bounded by one tile's duration and
// there is no cycle.  All CTAs must be co-resident for that argument: the grid is <= one CTA per SM, and launches that
// use the same `sync` words must be stream-ordered (the DiT step: one stream).
// Counters reset themselves: the last phase-1 consumer of a row block zeroes both words (kernel boundary = visibility).
// Results are bit-identical to the two-launch form: same tiles, same k-block order, same split boundaries.
#pragma once

namespace {

constexpr int CHAIN_MAX_ROW_BLOCKS = 64;                         // sync words: ready[64] | drained[64]
constexpr long long CHAIN_SYNC_BYTES = 2 * CHAIN_MAX_ROW_BLOCKS * 4;

struct ChainTiling {
#ifdef ECHO_DEBUG
  long long* trace;
#else
  static constexpr long long* trace = nullptr;
#endif
  int stages, stage_bytes;
  int tiles_m;                         // 256-row blocks (both phases)
  int tiles_n0, units0, kb_total0;     // phase 0: column tiles, units (= tiles_m * tiles_n0), k-blocks per unit
  int tiles_n1, splits1, units_mt1;    // phase 1: column tiles, split-K units per tile, units per row block
  int units1, kb_total1;
  FastDiv fd_n0, fd_u1, fd_s1, fd_m;   // divisions by tiles_n0 / units_mt1 / splits1 / tiles_m
  int m_fastest;                       // unit order inside a phase: 1 = row block fastest (pairs share a weight tile), 0 = column fastest
  long long split_stride1;
  int gate_sm;
  int ready_expect;                    // arrivals that complete a row block of h: 2 CTAs x EPI_WARPS x tiles_n0
  int drain_expect;                    // phase-1 A producers per row block: 2 CTAs x units_mt1
  int* ready;
  int* drained;
};

struct CUnit { int phase, mt, nt, split, kb0, kb1; };
struct ChainSched {
  int unit, step, end;
  __device__ ChainSched(const ChainTiling& T, int pair, int npairs) : unit(pair), step(npairs), end(T.units0 + T.units1) {}
  __device__ __forceinline__ bool next(const ChainTiling& T, CUnit& w) {
    if (unit >= end) return false;
    if (unit < T.units0) {
      if (T.m_fastest) { w.nt = T.fd_m.div(unit); w.mt = unit - w.nt * T.tiles_m; }
      else { w.mt = T.fd_n0.div(unit); w.nt = unit - w.mt * T.tiles_n0; }
      w.phase = 0; w.split = 0; w.kb0 = 0; w.kb1 = T.kb_total0;
    } else {
      const int v = unit - T.units0;
      int r;
      if (T.m_fastest) { r = T.fd_m.div(v); w.mt = v - r * T.tiles_m; }
      else { w.mt = T.fd_u1.div(v); r = v - w.mt * T.units_mt1; }
      w.phase = 1;
      w.nt = T.fd_s1.div(r); w.split = r - w.nt * T.splits1;
      w.kb0 = T.fd_s1.div(T.kb_total1 * w.split); w.kb1 = T.fd_s1.div(T.kb_total1 * (w.split + 1));
    }
    unit += step;
    return true;
  }
};

// A producer, first load of a phase-1 unit: every epilogue warp that writes rows of block `mt` of h has published
__device__ __forceinline__ void chain_wait_ready(const ChainTiling& T, int mt) {
  int v = 0;
  const long long t0 = clock64();
  do {
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(T.ready + mt) : "memory");
    if (v < T.ready_expect && clock64() - t0 > 8000000000LL) {
      printf("echo gemm chain: row block %d of the intermediate never completed (%d of %d)\n", mt, v, T.ready_expect);
      __trap();
    }
  } while (v < T.ready_expect);
  asm volatile("fence.proxy.async.global;" ::: "memory");      // the TMA (async proxy) reads below see those stores
  if (atomicAdd(T.drained + mt, 1) == T.drain_expect - 1) {    // last reader of the block: ready for the next launch
    *reinterpret_cast<volatile int*>(T.ready + mt) = 0;
    *reinterpret_cast<volatile int*>(T.drained + mt) = 0;
  }
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NTHREADS, 1)
gemm_chain_kernel(const __grid_constant__ CUtensorMap tmap_a0, const __grid_constant__ CUtensorMap tmap_b0,
                  const __grid_constant__ CUtensorMap tmap_a1, const __grid_constant__ CUtensorMap tmap_b1,
                  const __grid_constant__ CUtensorMap tmap_out1, const __grid_constant__ EpiParams P0,
                  const __grid_constant__ EpiParams P1, const __grid_constant__ ChainTiling T) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  float* scratch = reinterpret_cast<float*>(smem + RING_BYTES);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + RING_BYTES + SCRATCH_BYTES);
  uint64_t* empty_bar = full_bar + MAX_STAGES;
  uint64_t* tmem_full_bar = empty_bar + MAX_STAGES;
  uint64_t* tmem_empty_bar = tmem_full_bar + ACC_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + ACC_STAGES);
  float* gate_sm = reinterpret_cast<float*>(smem + RING_BYTES + SCRATCH_BYTES + 256);
  const int STAGES = T.stages, STAGE_SZ = T.stage_bytes;
  constexpr int BN = ACC_COLS, HALF_N = ACC_COLS / 2;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
#ifdef ECHO_DEBUG
  long long* const trace = T.trace ? T.trace + (long long)blockIdx.x * 32 : nullptr;
#else
  constexpr long long* trace = nullptr;
#endif
  if (trace && threadIdx.x == 0) trace[0] = clock64();
  ChainSched sch(T, pair, npairs);
  CUnit w;
  bool have = sch.next(T, w);

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_a0);
    ptx::prefetch_tensormap(&tmap_b0);
    ptx::prefetch_tensormap(&tmap_a1);
    ptx::prefetch_tensormap(&tmap_b1);
    ptx::prefetch_tensormap(&tmap_out1);
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < ACC_STAGES; ++a) {
      ptx::mbar_init(&tmem_full_bar[a], 1);
      ptx::mbar_init(&tmem_empty_bar[a], 2 * EPI_WARPS);
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
  pdl_launch_dependents();

  // do_a: the A operand (activations: xn in phase 0 -- after the grid dependency -- and h in phase 1 -- after its row
  // block is complete); do_b: the B operand (load-time weights: never waits)
  auto produce = [&](const bool do_a, const bool do_b) {
    int s = 0;
    uint32_t ph = 0;
    bool must_wait = do_a;
    long long waited = 0;
    for (; have; have = sch.next(T, w)) {
      const int m0 = w.mt * (2 * BM) + (int)rank * BM;
      const int n0 = w.nt * BN + (int)rank * HALF_N;
      const CUtensorMap* const ta = w.phase ? &tmap_a1 : &tmap_a0;
      const CUtensorMap* const tb = w.phase ? &tmap_b1 : &tmap_b0;
      bool need_ready = do_a && w.phase == 1;
      for (int kb = w.kb0; kb < w.kb1; ++kb) {
        ptx::mbar_wait(&empty_bar[s], ph ^ 1);
        uint8_t* sa = smem + s * STAGE_SZ;
        uint8_t* sb = sa + A_BYTES;
        if (do_a && rank == 0) ptx::mbar_arrive_expect_tx(&full_bar[s], 2u * (uint32_t)(A_BYTES + HALF_N * BK * 2));
        const uint32_t lbar = ptx::peer0_addr(&full_bar[s]);
        if (must_wait) {
          must_wait = false;
          pdl_wait();
          if (trace) trace[3] = clock64();
        }
        if (need_ready) {
          need_ready = false;
          const long long t0 = trace ? clock64() : 0;
          chain_wait_ready(T, w.mt);
          if (trace) waited += clock64() - t0;
        }
        if (do_a) ptx::tma_load_3d_cg2(sa, ta, lbar, kb * BK, m0, 0);
        if (do_b) ptx::tma_load_2d_cg2(sb, tb, lbar, kb * BK, n0);
        if (++s == STAGES) { s = 0; ph ^= 1; }
      }
    }
    if (trace && do_a) trace[4] = waited;     // clocks this CTA's A producer spent waiting for row blocks of h
  };
  if (warp == 0) {
    if (lane == 0) produce(true, false);
    __syncwarp();
  } else if (warp == B_PRODUCER_WARP) {
    if (lane == 0) produce(false, true);
    __syncwarp();
  } else if (warp == 1) {
    if (rank == 0) {
      const uint32_t idesc = ptx::umma_idesc_bf16(2 * BM, BN);
      uint32_t tc = 0, ph = 0;
      int s = 0;
      bool first = true;
      for (; have; have = sch.next(T, w), ++tc) {
        const int kb0 = w.kb0, kb1 = w.kb1;
        const uint32_t as = tc & 1, aph = (tc >> 1) & 1;
        if (lane == 0) ptx::mbar_wait(&tmem_empty_bar[as], aph ^ 1);
        __syncwarp();
        ptx::tc_fence_after();
        if (trace && tc < 6 && lane == 0) trace[8 + tc * 4 + 0] = clock64();
        const uint32_t tacc = tmem_base + as * ACC_COLS;
        for (int kb = kb0; kb < kb1; ++kb) {
          const uint32_t sa = smem_u32(smem + s * STAGE_SZ);
          const uint64_t da = ptx::umma_desc_sw128(sa);
          const uint64_t db = ptx::umma_desc_sw128(sa + A_BYTES);
          if (ptx::elect_one()) {
            ptx::mbar_wait(&full_bar[s], ph);
            ptx::tc_fence_after();
            if (trace && first) trace[2] = clock64();
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              ptx::umma_bf16_cg2(tacc, da + 2 * k, db + 2 * k, idesc, (kb != kb0 || k != 0) ? 1u : 0u);
            ptx::umma_commit_cg2_mc(&empty_bar[s], 0b11);
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
    pdl_wait();
    const int q = warp & 3;
    const int ew = warp - 2, half = ew >> 2;
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    uint32_t tc = 0;
    auto release_acc = [&](uint32_t as) {
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(ptx::peer0_addr(&tmem_empty_bar[as]));
    };
    for (; have; have = sch.next(T, w), ++tc) {
      const int m0 = w.mt * (2 * BM) + (int)rank * BM;
      const int nc0 = w.nt * BN;
      const uint32_t as = tc & 1, aph = (tc >> 1) & 1;
      TmemLoader2 L{tmem_base + as * ACC_COLS + lane_base};
      if (w.phase == 0) {
        ptx::mbar_wait(&tmem_full_bar[as], aph);
        ptx::tc_fence_after();
        if (trace && tc < 6 && warp == 2 && lane == 0) trace[8 + tc * 4 + 2] = clock64();
        EpiPre<ECHO_EPI_SWIGLU> pre;
        run_epilogue_t<ECHO_EPI_SWIGLU>(BN, P0, L, scratch + ew * EPI_SC_FLOATS, lane, half, 0, m0 + q * 32, nc0, pre, nullptr);
        release_acc(as);
        // this warp's rows x columns of h are written: publish (every lane fences its own stores, lane 0 counts the warp in)
        __threadfence();
        __syncwarp();
        if (lane == 0) asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(T.ready + w.mt), "r"(1) : "memory");
      } else {
        EpiPre<EPI_GENERIC_LEAN> pre;
        epilogue_prefetch_t<EPI_GENERIC_LEAN>(BN, P1, lane, half, 0, m0 + q * 32, nc0, pre);
        const float* gsm = nullptr;
        if (T.gate_sm && P1.gate && m0 < P1.rows) {     // the CTA's 128 rows share one gate row: stage it (see gemm_pair_kernel)
          const unsigned rpg = (unsigned)P1.rows_per_gate;
          const unsigned mg_first = (unsigned)m0, mg_last = mg_first + (unsigned)(min(BM, P1.rows - m0) - 1);
          const unsigned g_first = mg_first / rpg;
          if (g_first == mg_last / rpg) {
            asm volatile("bar.sync 3, %0;" ::"n"(32 * EPI_WARPS) : "memory");
            const int col = ew * 32 + lane;
            gate_sm[col] = nc0 + col < P1.n ? __ldg(P1.gate + (long long)g_first * P1.gate_ld + nc0 + col) : 0.f;
            asm volatile("bar.sync 3, %0;" ::"n"(32 * EPI_WARPS) : "memory");
            gsm = gate_sm;
          }
        }
        ptx::mbar_wait(&tmem_full_bar[as], aph);
        ptx::tc_fence_after();
        if (trace && tc < 6 && warp == 2 && lane == 0) trace[8 + tc * 4 + 2] = clock64();
        float* stage = scratch + ew * EPI_SC_FLOATS;
        int stage_stride = 0;
        if (sch.unit >= sch.end) {                      // the CTA's last unit: the idle operand ring stages every chunk
          stage = reinterpret_cast<float*>(smem) + ew * (ACC_COLS / 64) * EPI_SC_FLOATS;
          stage_stride = EPI_SC_FLOATS;
        }
        run_epilogue_t<EPI_GENERIC_LEAN>(BN, P1, L, stage, lane, half, 0, m0 + q * 32, nc0, pre, &tmap_out1, w.split,
                                         w.split * T.split_stride1, stage_stride, gsm);
        release_acc(as);
      }
      if (trace && tc < 6 && warp == 2 && lane == 0) trace[8 + tc * 4 + 3] = clock64();
    }
  }
  ptx::tc_fence_before();
  if (trace && threadIdx.x == 0) trace[6] = clock64();
  ptx::cluster_sync_all();
  if (warp == 1) ptx::tmem_dealloc_cg2(tmem_base, ACC_STAGES * ACC_COLS);
  if (trace && threadIdx.x == 0) trace[7] = clock64();
}

}  // namespace
