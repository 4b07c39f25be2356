// Joint attention for the DiT decoder and the causal speaker / latent-prefix encoders on tcgen05 (head_dim 128, up to four
// KV segments walked in place: self | latent prefix | text | speaker; a segment may be causal with a query offset).  Replaces torch.cat + F.scaled_dot_product_attention + the sigmoid
// gating of model.py:246-264 and the cache replication / scaling helpers of inference.py:409-428.
//
// One CTA = 128 queries of one (row, head).  Per 64-key tile:
//     S = Q K^T   tcgen05.mma  (A = Q from smem, B = K from smem, both K-major SW128; D = S in TMEM, 64 fp32 columns)
//     softmax     128 threads, one query row each: tcgen05.ld S -> registers, masks / extents / speaker_kv_scale,
//                 online max with LAZY rescale of O (only when the running max grows by > 8 in the exp2 domain),
//                 P = exp2(s - m) as bf16 written back to TMEM over the first 32 columns of S (tcgen05.st)
//     O += P V    tcgen05.mma  (A = P from TMEM, B = V from smem as an MN-major SW128 operand: no transpose copy)
// S is double-buffered in TMEM so the QK^T of tile j+1 is issued while the softmax of tile j runs; K and V travel
// in separate 2-stage TMA rings (K is needed one iteration earlier than V).  A CTA needs 96 KB of shared memory
// and 256 TMEM columns, so two CTAs are co-resident per SM and fill each other's bubbles.  Warp roles: 0..3 softmax
// + epilogue (sigmoid gate, 1/l, bf16 store), 4 = TMA producer, 5 = TMEM owner + MMA issuer.
#include "common.cuh"
#include "ptx_sm100.cuh"
#include "../../include/echo_b200.h"

int echo_make_tmap_bf16(CUtensorMap* m, const void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides_b,
                        const cuuint32_t* box);   // gemm.cu

namespace {
constexpr int QT = 128, KT = 64, HD = 128;
constexpr int Q_BYTES = QT * HD * 2;          // 32 KB: two [128 x 64] SW128 blocks
constexpr int KV_BYTES = KT * HD * 2;         // 16 KB: two [64 x 64] SW128 blocks
constexpr int MAX_TILES = 128;                // <= 8192 keys in total
constexpr int SMEM_ATTN = Q_BYTES + 4 * KV_BYTES + 1024 /*align*/ + 256 /*barriers*/ + MAX_TILES * 16;
constexpr int TM_S0 = 0, TM_S1 = 64, TM_O = 128, TM_COLS = 256;
constexpr float RESCALE_TAU = 8.f;            // lazy-rescale threshold (log2 domain): P <= 2^8
// split-KV exchange: one part = un-normalised O [32 float4 column groups][128 rows] (a warp reads / writes 512
// contiguous bytes per instruction) + m[128] + l[128]
constexpr int PART_FLOATS = QT * HD + 2 * QT;
constexpr int WS_COUNTERS = 16384;            // ints at the head of the workspace
constexpr int MAX_SPLITS = 4;

struct AttnSeg {
  int len, kv_batch_mod;
  int causal, q_offset;  // causal: key j visible to query i iff j <= i + q_offset (tiles past the diagonal are never
                         // listed; tiles ON it carry a flag and are masked per row)
  const uint8_t* mask; long long mask_stride;
  const int* extent;
  const float* kv_scale;
};
struct AttnParams {
#ifdef ECHO_DEBUG
  long long* trace;      // nullptr unless tracing: [cta][64] clock64 stamps (debug builds only)
#else
  static constexpr long long* trace = nullptr;
#endif
  int rows, seq, heads, n_seg;
  int splits;            // > 1: split-KV -- `splits` CTAs per (row, head, query tile), each a contiguous share of the key
                         // tiles; partial (O, m, l) through `ws_parts`, every part merges its share of the columns
  float* ws_parts;       // [group][split][PART_FLOATS]
  int* ws_count;         // [group][2]: parts arrived / parts done merging; zero between launches (the last part resets)
  int static_first;      // segment 0 has no mask / extent and len > 0: tile 0 is known before the tile list is built
  float softmax_scale;
  const bf16* gate;
  bf16* out;
  AttnSeg seg[4];
};
struct TileInfo { int seg, key0; unsigned long long valid; };   // 16 bytes; seg bit 8: diagonal tile of a causal segment

__device__ __forceinline__ void tma_load_4d(void* dst, const void* tmap, uint64_t* bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
// MN-major SW128 operand ([k rows][64 elements = 128 B] blocks; 8-row groups 1024 B apart; next 64-column block `lbo` B on)
__device__ __forceinline__ uint64_t umma_desc_sw128_mn(uint32_t smem_addr, uint32_t lbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)(lbo_bytes >> 4) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// D[tmem] (+)= A[tmem] * B[smem]
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld32_nowait(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ float ex2_approx(float x) {   // 2^x, -inf -> 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
        "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
        "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
        "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tma_store_4d(const void* tmap, const void* src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit_wait() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

struct Maps { CUtensorMap q, k[4], v[4], gate, out; };
#ifdef ECHO_DEBUG
static long long* g_attn_trace = nullptr;   // debug timeline, see echo_debug_attn_trace
#endif

__global__ void __launch_bounds__(192, 2) attention_tc_kernel(const __grid_constant__ Maps M,
                                                              const __grid_constant__ AttnParams A) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + Q_BYTES;                     // 2 stages
  uint8_t* sV = sK + 2 * KV_BYTES;                  // 2 stages
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + 2 * KV_BYTES);
  uint64_t* q_full = bars;            // 1
  uint64_t* k_full = bars + 1;        // 2
  uint64_t* k_empty = bars + 3;       // 2
  uint64_t* v_full = bars + 5;        // 2
  uint64_t* v_empty = bars + 7;       // 2
  uint64_t* s_full = bars + 9;        // 2
  uint64_t* p_full = bars + 11;       // 2
  uint64_t* pv_done = bars + 13;      // 2
  uint64_t* g_full = bars + 15;       // 1: gate tile landed in the (drained) K ring
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);
  int* s_ntiles = reinterpret_cast<int*>(bars + 17);
  TileInfo* tiles = reinterpret_cast<TileInfo*>(reinterpret_cast<uint8_t*>(bars) + 256);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nsplit = A.splits > 1 ? A.splits : 1;
  const int qt = (int)blockIdx.x / nsplit, part = (int)blockIdx.x - qt * nsplit;
  const int q0 = qt * QT, head = blockIdx.y, row = blockIdx.z;
#ifdef ECHO_DEBUG
  long long* const trace = A.trace ? A.trace + ((long long)(blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 64 : nullptr;
#else
  constexpr long long* trace = nullptr;
#endif
  if (trace && tid == 0) {
    trace[0] = clock64();
    unsigned long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    trace[57] = (long long)gt;                       // ns, comparable across SMs
  }
  pdl_launch_dependents();

  // ---- tile list: every 64-key tile any query of this row can see, with its key-valid bits.  Two memory round
  //      trips for the whole CTA: all extents, then one tile per lane with 16-byte mask loads. ---------------------
  if (warp == 0) {
    pdl_wait();            // extents / masks may come from the previous kernel
    if (trace && tid == 0) trace[2] = clock64();
    // extents and mask bytes are loaded together: lanes walk the POTENTIAL tiles (from the static segment lengths),
    // then tiles with no visible key are squeezed out with a ballot (order preserved)
    int ext_l = 0x7fffffff;
    if (lane < A.n_seg && A.seg[lane].extent) ext_l = A.seg[lane].extent[row];
    int pbase[5];
    pbase[0] = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) pbase[i + 1] = pbase[i] + (i < A.n_seg ? (max(A.seg[i].len, 0) + KT - 1) / KT : 0);
    int nkept = 0;
    if (trace && tid == 0) trace[56] = clock64() + (ext_l & 1) - (ext_l & 1);
    for (int g0 = 0; g0 < pbase[4]; g0 += 32) {
      const int g = g0 + lane;
      const bool in = g < pbase[4];
      const int sgi = in ? (g >= pbase[1]) + (g >= pbase[2]) + (g >= pbase[3]) : 0;
      const int b0 = sgi == 0 ? 0 : (sgi == 1 ? pbase[1] : (sgi == 2 ? pbase[2] : pbase[3]));
      const AttnSeg& sg = A.seg[sgi];
      const int key0 = (g - b0) * KT;
      unsigned long long mbits = ~0ull;
      if (in && sg.mask) {
        // 64 mask bytes -> 64 bits.  One path for aligned and ragged rows (the 53-key speaker segment): 17 aligned
        // 32-bit loads, all in flight together, realigned with funnel shifts; words wholly outside the row are not
        // read (an aligned word holding one valid byte cannot leave the allocation).
        const uint8_t* mrow = sg.mask + (long long)row * sg.mask_stride + key0;
        const int nvalid = min(sg.len - key0, 64);
        const unsigned mis = (unsigned)(reinterpret_cast<uintptr_t>(mrow) & 3);
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(mrow - mis);
        uint32_t w[17];
#pragma unroll
        for (int i = 0; i < 17; ++i) w[i] = (i * 4 < (int)mis + nvalid) ? wp[i] : 0u;
        uint32_t bits[2] = {0u, 0u};
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const uint32_t v = __funnelshift_r(w[i], w[i + 1], mis * 8);
          const uint32_t nz = (((v & 0x7f7f7f7fu) + 0x7f7f7f7fu) | v) & 0x80808080u;   // bit 7 of every non-zero byte
          bits[i >> 3] |= ((((nz >> 7) * 0x01020408u) >> 24) & 0xfu) << ((i & 7) * 4);
        }
        mbits = ((unsigned long long)bits[1] << 32) | bits[0];
        if (nvalid < 64) mbits &= (1ull << nvalid) - 1ull;
      }
      int e = min(__shfl_sync(0xffffffffu, ext_l, sgi), sg.len);
      int diag = 0;
      if (in && sg.causal) {                               // keys past the last query's horizon are never visited
        e = min(e, q0 + QT + sg.q_offset);                 // key j <= (q0 + QT - 1) + q_offset
        diag = key0 + KT - 1 > q0 + sg.q_offset;           // the tile's last key is hidden from the tile's first query
      }
      const unsigned long long valid =
          (!in || e <= key0) ? 0ull : (mbits & ((e - key0 >= 64) ? ~0ull : ((1ull << (e - key0)) - 1ull)));
      const unsigned keep = __ballot_sync(0xffffffffu, valid != 0ull);
      if (valid != 0ull) {
        const int pos = nkept + __popc(keep & ((1u << lane) - 1u));
        tiles[pos].seg = sgi | (diag << 8); tiles[pos].key0 = key0; tiles[pos].valid = valid;
      }
      nkept += __popc(keep);
    }
    if (lane == 0) *s_ntiles = nkept;
    if (trace && tid == 0) trace[3] = clock64();
  }
  if (warp == 4 && lane == 0) {
    ptx::prefetch_tensormap(&M.q);
    ptx::prefetch_tensormap(&M.k[0]);
    ptx::prefetch_tensormap(&M.v[0]);
    ptx::mbar_init(q_full, 1);
    ptx::mbar_init(g_full, 1);
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&k_full[i], 1); ptx::mbar_init(&k_empty[i], 1);
      ptx::mbar_init(&v_full[i], 1); ptx::mbar_init(&v_empty[i], 1);
      ptx::mbar_init(&s_full[i], 1); ptx::mbar_init(&p_full[i], 128); ptx::mbar_init(&pv_done[i], 1);
    }
    ptx::fence_barrier_init();
    if (A.static_first && part == 0) {
      // Q and the first K / V tile do not depend on the tile list: put them in flight under the rest of the set-up.
      pdl_wait();
      const int kvrow = A.seg[0].kv_batch_mod > 0 ? row % A.seg[0].kv_batch_mod : row;
      ptx::mbar_arrive_expect_tx(q_full, Q_BYTES);
      tma_load_4d(sQ, &M.q, q_full, 0, head, q0, row);
      tma_load_4d(sQ + Q_BYTES / 2, &M.q, q_full, 64, head, q0, row);
      ptx::mbar_arrive_expect_tx(&k_full[0], KV_BYTES);
      tma_load_4d(sK, &M.k[0], &k_full[0], 0, head, 0, kvrow);
      tma_load_4d(sK + KV_BYTES / 2, &M.k[0], &k_full[0], 64, head, 0, kvrow);
      ptx::mbar_arrive_expect_tx(&v_full[0], KV_BYTES);
      tma_load_4d(sV, &M.v[0], &v_full[0], 0, head, 0, kvrow);
      tma_load_4d(sV + KV_BYTES / 2, &M.v[0], &v_full[0], 64, head, 0, kvrow);
    }
  }
  if (warp == 5) {
    ptx::tmem_alloc(tmem_slot, TM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // split-KV: this CTA's contiguous share of the tile list (every part builds the same list)
  const int ntiles_all = *s_ntiles;
  const int nparts = ntiles_all < nsplit ? (ntiles_all > 0 ? ntiles_all : 1) : nsplit;     // parts that have tiles
  const int j_first = part < nparts ? (int)((long long)ntiles_all * part / nparts) : 0;
  const int ntiles = part < nparts ? (int)((long long)ntiles_all * (part + 1) / nparts) - j_first : 0;
  tiles += j_first;
  const bool idle_part = part >= nparts;      // more parts than tiles: nothing to do, nothing to publish
  const bool merging = nparts > 1;            // split-KV: output written by the parts' merge, not by the TMA epilogue
  pdl_wait();
  if (trace && tid == 0) trace[1] = clock64();

  if (warp == 4) {
    // ===== TMA producer =====
    if (lane == 0 && ntiles > 0) {
      const bool early = A.static_first != 0 && part == 0;     // Q, K(0), V(0) already in flight
      if (!early) {
        ptx::mbar_arrive_expect_tx(q_full, Q_BYTES);
        tma_load_4d(sQ, &M.q, q_full, 0, head, q0, row);
        tma_load_4d(sQ + Q_BYTES / 2, &M.q, q_full, 64, head, q0, row);
      }
      auto load_k = [&](int j) {
        const int s = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        TileInfo t = tiles[j];
        t.seg &= 0xff;
        const int kvrow = A.seg[t.seg].kv_batch_mod > 0 ? row % A.seg[t.seg].kv_batch_mod : row;
        ptx::mbar_wait(&k_empty[s], ph ^ 1);
        ptx::mbar_arrive_expect_tx(&k_full[s], KV_BYTES);
        tma_load_4d(sK + s * KV_BYTES, &M.k[t.seg], &k_full[s], 0, head, t.key0, kvrow);
        tma_load_4d(sK + s * KV_BYTES + KV_BYTES / 2, &M.k[t.seg], &k_full[s], 64, head, t.key0, kvrow);
      };
      auto load_v = [&](int j) {
        const int s = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        TileInfo t = tiles[j];
        t.seg &= 0xff;
        const int kvrow = A.seg[t.seg].kv_batch_mod > 0 ? row % A.seg[t.seg].kv_batch_mod : row;
        ptx::mbar_wait(&v_empty[s], ph ^ 1);
        ptx::mbar_arrive_expect_tx(&v_full[s], KV_BYTES);
        tma_load_4d(sV + s * KV_BYTES, &M.v[t.seg], &v_full[s], 0, head, t.key0, kvrow);
        tma_load_4d(sV + s * KV_BYTES + KV_BYTES / 2, &M.v[t.seg], &v_full[s], 64, head, t.key0, kvrow);
      };
      if (!early) load_k(0);
      for (int j = 0; j < ntiles; ++j) {      // K runs one tile ahead of V
        if (j + 1 < ntiles) load_k(j + 1);
        if (j > 0 || !early) load_v(j);
      }
      if (A.gate && !merging) {
        // every QK^T has been issued: once the last one per stage retires, the K ring (32 KB) is free to take this
        // CTA's [128 x 128] gate tile, which then doubles as the staging buffer of the output TMA store
        const int jl = ntiles - 1;
        ptx::mbar_wait(&k_empty[jl & 1], (jl >> 1) & 1);
        if (jl > 0) ptx::mbar_wait(&k_empty[(jl - 1) & 1], ((jl - 1) >> 1) & 1);
        ptx::mbar_arrive_expect_tx(g_full, Q_BYTES);
        tma_load_4d(sK, &M.gate, g_full, 0, head, q0, row);
        tma_load_4d(sK + Q_BYTES / 2, &M.gate, g_full, 64, head, q0, row);
      }
    }
    __syncwarp();
  } else if (warp == 5) {
    // ===== MMA issuer: whole warp converged, one elected lane waits and issues =====
    if (ntiles > 0) {
      constexpr uint32_t idesc_s = ptx::umma_idesc_bf16(QT, KT);                    // 128 x 64, both K-major
      constexpr uint32_t idesc_o = ptx::umma_idesc_bf16(QT, HD) | (1u << 16);       // 128 x 128, B (= V) MN-major
      const uint32_t q_addr = smem_u32(sQ);
      auto issue_s = [&](int j) {
        const int s = j & 1;
        const uint32_t k_addr = smem_u32(sK + s * KV_BYTES);
        const uint32_t d = tmem_base + (s ? TM_S1 : TM_S0);
        if (ptx::elect_one()) {
          ptx::mbar_wait(&k_full[s], (j >> 1) & 1);
          ptx::tc_fence_after();
#pragma unroll
          for (int kk = 0; kk < HD / 16; ++kk) {
            const uint64_t da = ptx::umma_desc_sw128(q_addr + (kk >> 2) * (Q_BYTES / 2)) + 2 * (kk & 3);
            const uint64_t db = ptx::umma_desc_sw128(k_addr + (kk >> 2) * (KV_BYTES / 2)) + 2 * (kk & 3);
            ptx::umma_bf16(d, da, db, idesc_s, kk != 0 ? 1u : 0u);
          }
          ptx::umma_commit(&k_empty[s]);
          ptx::umma_commit(&s_full[s]);
        }
        __syncwarp();
      };
      if (ptx::elect_one()) {
        if (trace) trace[52] = clock64();
        ptx::mbar_wait(q_full, 0);
        if (trace) trace[53] = clock64();
        ptx::mbar_wait(&k_full[0], 0);
        if (trace) trace[54] = clock64();
      }
      __syncwarp();
      issue_s(0);
      if (trace && ptx::elect_one()) trace[55] = clock64();
      __syncwarp();
      for (int j = 0; j < ntiles; ++j) {
        const int s = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        if (j + 1 < ntiles) issue_s(j + 1);
        const uint32_t v_addr = smem_u32(sV + s * KV_BYTES);
        const uint32_t p_tmem = tmem_base + (s ? TM_S1 : TM_S0);
        if (ptx::elect_one()) {
          ptx::mbar_wait(&p_full[s], ph);
          ptx::mbar_wait(&v_full[s], ph);
          ptx::tc_fence_after();
#pragma unroll
          for (int kk = 0; kk < KT / 16; ++kk) {
            const uint64_t db = umma_desc_sw128_mn(v_addr + kk * 16 * 128, KV_BYTES / 2);
            umma_bf16_ts(tmem_base + TM_O, p_tmem + kk * 8, db, idesc_o, (j | kk) != 0 ? 1u : 0u);
          }
          ptx::umma_commit(&v_empty[s]);
          ptx::umma_commit(&pv_done[s]);
          if (trace && j < 12) trace[4 + 4 * j + 3] = clock64();
        }
        __syncwarp();
      }
    }
    __syncwarp();
  } else {
    // ===== softmax + epilogue: thread = query row =====
    const uint32_t lane_off = (uint32_t)(warp * 32) << 16;
    const float sl2 = A.softmax_scale * 1.4426950408889634f;
    float m = -INFINITY, l = 0.f;
    float fseg[4];                                         // per-segment K/V multiplier (speaker_kv_scale), loaded once
#pragma unroll
    for (int i = 0; i < 4; ++i) fseg[i] = (i < A.n_seg && A.seg[i].kv_scale) ? __ldg(A.seg[i].kv_scale) : 1.f;
    for (int j = 0; j < ntiles; ++j) {
      const int b = j & 1;
      const uint32_t ph = (j >> 1) & 1;
      TileInfo t = tiles[j];
      const bool diag = (t.seg >> 8) != 0;
      t.seg &= 0xff;
      const float f = t.seg == 0 ? fseg[0] : (t.seg == 1 ? fseg[1] : (t.seg == 2 ? fseg[2] : fseg[3]));
      const float sc = sl2 * f;
      ptx::mbar_wait(&s_full[b], ph);
      ptx::tc_fence_after();
      if (trace && tid == 0 && j < 12) trace[4 + 4 * j + 0] = clock64();
      float s[64];
      const uint32_t s_addr = tmem_base + (b ? TM_S1 : TM_S0) + lane_off;
      tmem_ld32_nowait(s_addr, s);
      tmem_ld32_nowait(s_addr + 32, s + 32);
      tmem_ld_wait();
      if (t.valid != ~0ull || diag) {                      // partial tile (warp-uniform test; most tiles are full)
        unsigned long long vm = t.valid;
        if (diag) {                                        // causal diagonal: this row sees keys <= q0 + tid + q_offset
          const int nvis = q0 + tid + A.seg[t.seg].q_offset - t.key0 + 1;
          vm &= nvis >= 64 ? ~0ull : (nvis <= 0 ? 0ull : ((1ull << nvis) - 1ull));
        }
        const unsigned lo = (unsigned)vm, hi = (unsigned)(vm >> 32);
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          if (!((lo >> c) & 1u)) s[c] = -INFINITY;
          if (!((hi >> c) & 1u)) s[c + 32] = -INFINITY;
        }
      }
      uint32_t p[32];
      float lsum = 0.f, mt;
      // P = 2^(s*sc - m) [* f], its row sum and the raw row maximum in ONE pass (the max instructions hide under the
      // MUFU-bound exponentials)
      auto pass = [&](float nm, bool scaled) {
        float x0 = -INFINITY, x1 = -INFINITY, x2 = -INFINITY, x3 = -INFINITY, l0 = 0.f, l1 = 0.f;
#pragma unroll
        for (int c = 0; c < 64; c += 4) {
          x0 = fmaxf(x0, s[c]); x1 = fmaxf(x1, s[c + 1]); x2 = fmaxf(x2, s[c + 2]); x3 = fmaxf(x3, s[c + 3]);
          const float p0 = ex2_approx(fmaf(s[c], sc, nm)), p1 = ex2_approx(fmaf(s[c + 1], sc, nm));
          const float p2 = ex2_approx(fmaf(s[c + 2], sc, nm)), p3 = ex2_approx(fmaf(s[c + 3], sc, nm));
          l0 += p0 + p1; l1 += p2 + p3;
          p[c >> 1] = scaled ? pack_bf16x2(p0 * f, p1 * f) : pack_bf16x2(p0, p1);
          p[(c >> 1) + 1] = scaled ? pack_bf16x2(p2 * f, p3 * f) : pack_bf16x2(p2, p3);
        }
        lsum = l0 + l1;
        mt = fmaxf(fmaxf(x0, x1), fmaxf(x2, x3)) * sc;
      };
      // Speculate that the running maximum still covers this tile (true for almost every tile after the first): the
      // exponentials then do not wait for the row maximum.  Rows that have not seen a valid key yet (m = -inf) and
      // tiles whose maximum outgrows m by more than 2^8 take the exact path below and redo the pass.
      const bool fresh = __any_sync(0xffffffffu, m == -INFINITY);
      if (!fresh) {
        if (f == 1.f) pass(-m, false); else pass(-m, true);
      } else {
        float x0 = -INFINITY, x1 = -INFINITY, x2 = -INFINITY, x3 = -INFINITY;
#pragma unroll
        for (int c = 0; c < 64; c += 4) {
          x0 = fmaxf(x0, s[c]); x1 = fmaxf(x1, s[c + 1]); x2 = fmaxf(x2, s[c + 2]); x3 = fmaxf(x3, s[c + 3]);
        }
        mt = fmaxf(fmaxf(x0, x1), fmaxf(x2, x3)) * sc;
      }
      const bool grow = mt > m + RESCALE_TAU;            // also true for the first finite max (m = -inf)
      if (fresh || __any_sync(0xffffffffu, grow)) {
        // the PV of the previous tile must have landed before O may be rescaled
        if (j > 0) ptx::mbar_wait(&pv_done[(j - 1) & 1], ((j - 1) >> 1) & 1);
        const float mnew = grow ? mt : m;
        const float alpha = (m == -INFINITY) ? 0.f : exp2f(m - mnew);
        m = mnew;
        l *= alpha;
        if (j > 0) {
          ptx::tc_fence_after();
#pragma unroll 1
          for (int c = 0; c < HD; c += 32) {
            float o[32];
            ptx::tmem_ld32(tmem_base + TM_O + lane_off + c, o);
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] *= alpha;
            tmem_st32(tmem_base + TM_O + lane_off + c, reinterpret_cast<const uint32_t*>(o));
          }
          tmem_st_wait();
        }
        pass((m == -INFINITY) ? 0.f : -m, true);
      }
      if (trace && tid == 0 && j < 12) trace[4 + 4 * j + 1] = clock64();
      l += lsum;
      tmem_st32(s_addr, p);
      tmem_st_wait();
      ptx::tc_fence_before();
      ptx::mbar_arrive(&p_full[b]);
      if (trace && tid == 0 && j < 12) trace[4 + 4 * j + 2] = clock64();
    }
    // ---- epilogue: O / l * sigmoid-gate -> bf16, staged in the K ring (SW128 rows: conflict-free 16-byte accesses
    //      for a thread-per-row pattern) and written with two TMA stores (rows past `seq` are clipped by the map) ----
    if (ntiles > 0) {
      ptx::mbar_wait(&pv_done[(ntiles - 1) & 1], ((ntiles - 1) >> 1) & 1);
      ptx::tc_fence_after();
      if (A.gate && !merging) ptx::mbar_wait(g_full, 0);
    }
    if (trace && tid == 0) trace[60] = clock64();
    // ---- split-KV: every part publishes its un-normalised (O, m, l), waits until all parts of the group have (they are
    //      co-resident: the host splits only when groups x parts fit the SMs' CTA slots), then merges ITS share of the
    //      128 output columns (32-column chunks dealt round-robin) over all parts in part order and writes it: the merge
    //      traffic is spread over the group's SMs instead of 64 KB x parts into one, and the result does not depend on
    //      arrival order.  The last part through resets the group's two counters for the next launch. ----
    bool do_out = !idle_part;
    if (merging && !idle_part) {
      const int group = ((int)row * A.heads + head) * ((int)gridDim.x / nsplit) + qt;
      const float* parts = A.ws_parts + (long long)group * nsplit * PART_FLOATS;
      float* mine = A.ws_parts + ((long long)group * nsplit + part) * PART_FLOATS;
      int* cnt = A.ws_count + 2 * group;
#pragma unroll 1
      for (int c = 0; c < HD; c += 32) {
        float o[32];
        ptx::tmem_ld32(tmem_base + TM_O + lane_off + c, o);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          __stcg(reinterpret_cast<float4*>(mine) + ((c >> 2) + k) * QT + tid,
                 make_float4(o[4 * k], o[4 * k + 1], o[4 * k + 2], o[4 * k + 3]));
      }
      __stcg(mine + QT * HD + tid, m);
      __stcg(mine + QT * HD + QT + tid, l);
      if (trace && tid == 0) trace[48] = clock64();
      __threadfence();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (trace && tid == 0) trace[49] = clock64();
      if (tid == 0) {
        atomicAdd(cnt, 1);
        int v = 0;
        const long long t0 = clock64();
        for (;;) {
          asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(cnt) : "memory");
          if (v >= nparts) break;
          if (clock64() - t0 > 4000000000LL) { printf("echo attention: split-KV part never arrived\n"); __trap(); }
          __nanosleep(32);
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (trace && tid == 0) trace[50] = clock64();
      // (the partials were released at gpu scope before each part's count; they are read from the L2 below: no L1 copy
      // can be stale and no second fence is needed)
      // Everything the merge needs is requested before anything is used: the parts' (m, l), the gate values and the
      // first two parts' columns share ONE memory round trip; later parts follow two at a time.
      float mp[MAX_SPLITS], lp[MAX_SPLITS];
#pragma unroll
      for (int p2 = 0; p2 < MAX_SPLITS; ++p2) {
        mp[p2] = p2 < nparts ? __ldcg(parts + (long long)p2 * PART_FLOATS + QT * HD + tid) : -INFINITY;
        lp[p2] = p2 < nparts ? __ldcg(parts + (long long)p2 * PART_FLOATS + QT * HD + QT + tid) : 0.f;
      }
      const bool row_ok = q0 + tid < A.seq;
      const long long oidx = ((long long)row * A.seq + q0 + tid) * ((long long)A.heads * HD) + (long long)head * HD;
      float wp[MAX_SPLITS], invm = 0.f;
#pragma unroll 1
      for (int c = 32 * part; c < HD; c += 32 * nparts) {
        uint4 gv[4];
        if (A.gate && row_ok) {
#pragma unroll
          for (int g = 0; g < 4; ++g) gv[g] = __ldg(reinterpret_cast<const uint4*>(A.gate + oidx + c + 8 * g));
        }
        float4 t0[8], t1[8];
        const float4* src = reinterpret_cast<const float4*>(parts) + (c >> 2) * QT + tid;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          t0[k] = __ldcg(src + k * QT);
          t1[k] = __ldcg(src + (PART_FLOATS >> 2) + k * QT);                  // nparts >= 2 here
        }
        if (c == 32 * part) {                                                // weights of the parts, once
          float mx = -INFINITY;
#pragma unroll
          for (int p2 = 0; p2 < MAX_SPLITS; ++p2) mx = fmaxf(mx, mp[p2]);
          float lm = 0.f;
#pragma unroll
          for (int p2 = 0; p2 < MAX_SPLITS; ++p2) {
            wp[p2] = (p2 < nparts && mp[p2] != -INFINITY) ? exp2f(mp[p2] - mx) : 0.f;
            lm = fmaf(wp[p2], lp[p2], lm);
          }
          invm = lm > 0.f ? 1.f / lm : 0.f;
        }
        float o[32];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          o[4 * k] = fmaf(wp[1], t1[k].x, wp[0] * t0[k].x); o[4 * k + 1] = fmaf(wp[1], t1[k].y, wp[0] * t0[k].y);
          o[4 * k + 2] = fmaf(wp[1], t1[k].z, wp[0] * t0[k].z); o[4 * k + 3] = fmaf(wp[1], t1[k].w, wp[0] * t0[k].w);
        }
        if (nparts > 2) {
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            t0[k] = __ldcg(src + 2 * (PART_FLOATS >> 2) + k * QT);
            t1[k] = nparts > 3 ? __ldcg(src + 3 * (PART_FLOATS >> 2) + k * QT) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            o[4 * k] = fmaf(wp[3], t1[k].x, fmaf(wp[2], t0[k].x, o[4 * k]));
            o[4 * k + 1] = fmaf(wp[3], t1[k].y, fmaf(wp[2], t0[k].y, o[4 * k + 1]));
            o[4 * k + 2] = fmaf(wp[3], t1[k].z, fmaf(wp[2], t0[k].z, o[4 * k + 2]));
            o[4 * k + 3] = fmaf(wp[3], t1[k].w, fmaf(wp[2], t0[k].w, o[4 * k + 3]));
          }
        }
        if (row_ok) {
#pragma unroll
          for (int g = 0; g < 32; g += 8) {
            float w[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) w[i] = o[g + i] * invm;
            if (A.gate) {
              const uint4 gvv = gv[g >> 3];
              const uint4 gv = gvv;
              const float2 g0 = unpack_bf16x2(gv.x), g1 = unpack_bf16x2(gv.y), g2 = unpack_bf16x2(gv.z), g3 = unpack_bf16x2(gv.w);
              w[0] *= g0.x; w[1] *= g0.y; w[2] *= g1.x; w[3] *= g1.y; w[4] *= g2.x; w[5] *= g2.y; w[6] *= g3.x; w[7] *= g3.y;
            }
            uint4 u;
            u.x = pack_bf16x2(w[0], w[1]); u.y = pack_bf16x2(w[2], w[3]);
            u.z = pack_bf16x2(w[4], w[5]); u.w = pack_bf16x2(w[6], w[7]);
            *reinterpret_cast<uint4*>(A.out + oidx + c + g) = u;
          }
        }
      }
      if (trace && tid == 0) trace[51] = clock64();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (tid == 0 && atomicAdd(cnt + 1, 1) == nparts - 1) { cnt[0] = 0; cnt[1] = 0; }
      do_out = false;
    }
    const float inv = l > 0.f ? 1.f / l : 0.f;
    uint8_t* const stage_row = sK + tid * 128;
#pragma unroll 1
    for (int c = 0; c < HD && do_out; c += 32) {
      float o[32];
      if (ntiles > 0) {
        ptx::tmem_ld32(tmem_base + TM_O + lane_off + c, o);
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = 0.f;
      }
#pragma unroll
      for (int g = 0; g < 32; g += 8) {
        const int col = c + g;
        uint4* cell = reinterpret_cast<uint4*>(stage_row + (col >> 6) * (Q_BYTES / 2) + ((((col & 63) >> 3) ^ (tid & 7)) << 4));
        float w[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) w[i] = o[g + i] * inv;
        if (A.gate && ntiles > 0) {
          const uint4 gv = *cell;
          const float2 g0 = unpack_bf16x2(gv.x), g1 = unpack_bf16x2(gv.y), g2 = unpack_bf16x2(gv.z), g3 = unpack_bf16x2(gv.w);
          w[0] *= g0.x; w[1] *= g0.y; w[2] *= g1.x; w[3] *= g1.y; w[4] *= g2.x; w[5] *= g2.y; w[6] *= g3.x; w[7] *= g3.y;
        }
        uint4 u;
        u.x = pack_bf16x2(w[0], w[1]); u.y = pack_bf16x2(w[2], w[3]);
        u.z = pack_bf16x2(w[4], w[5]); u.w = pack_bf16x2(w[6], w[7]);
        *cell = u;
      }
    }
    ptx::fence_proxy_async();                              // generic-proxy writes -> visible to the TMA store
    asm volatile("bar.sync 1, 128;" ::: "memory");         // the four softmax warps only
    if (tid == 0 && do_out) {
      tma_store_4d(&M.out, sK, 0, head, q0, row);
      tma_store_4d(&M.out, sK + Q_BYTES / 2, 64, head, q0, row);
      tma_store_commit_wait();
    }
  }
  if (trace && tid == 0) trace[61] = clock64();
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 5) ptx::tmem_dealloc(tmem_base, TM_COLS);
  if (trace && tid == 0) {
    trace[62] = clock64(); trace[63] = ntiles;
    unsigned long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    trace[58] = (long long)gt;
  }
}
}  // namespace

// Bytes of the split-KV exchange buffer (echo_attn_desc.workspace) that lets a call with this many (row, head, query
// tile) groups use up to `kv_splits` parts: counters + one partial (O, m, l) per part.  Zeroed once by the caller.
extern "C" long long echo_attention_workspace_bytes(int rows, int seq, int heads, int kv_splits) {
  if (kv_splits < 1) kv_splits = MAX_SPLITS;
  if (kv_splits > MAX_SPLITS) kv_splits = MAX_SPLITS;
  const long long groups = (long long)ceil_div(seq, QT) * heads * rows;
  return (long long)WS_COUNTERS * 4 + groups * kv_splits * (long long)PART_FLOATS * 4;
}

// Debug aid: >= 64 int64 per CTA; [0] start, [1] setup done, per tile j < 12: [4+4j] S(j) ready, [5+4j] PV(j-1) done,
// [6+4j] P(j) published, [7+4j] PV(j) issued (MMA lane); [2] warp 0 past the PDL wait, [3] tile list built,
// [60] last PV done (+ gate tile landed), [61] epilogue done, [62] exit, [63] #tiles.
extern "C" int echo_debug_attn_trace(long long* device_buf) {
#ifdef ECHO_DEBUG
  g_attn_trace = device_buf;
  return ECHO_OK;
#else
  if (!device_buf) return ECHO_OK;
  echo_set_error("echo_debug_attn_trace: this library was built without -DECHO_DEBUG (make -C echo_tts_b200/csrc DEBUG=1)");
  return ECHO_ERR_UNSUPPORTED;
#endif
}

// Returns ECHO_ERR_UNSUPPORTED (without setting an error) when the descriptor is outside this kernel's domain so
// that echo_attention() can fall back to the generic mma.sync kernel (causal / windowed / head_dim 64 encoders).
int echo_attention_tc(const echo_attn_desc* d, cudaStream_t st) {
  if (d->head_dim != HD) return ECHO_ERR_UNSUPPORTED;
  long long total_tiles = 0;
  for (int i = 0; i < d->n_seg; ++i) {
    if (d->seg[i].window > 0) return ECHO_ERR_UNSUPPORTED;
    if (d->seg[i].causal && d->seg[i].q_offset < 0) return ECHO_ERR_UNSUPPORTED;
    total_tiles += ceil_div(d->seg[i].len, KT);
  }
  if (total_tiles > MAX_TILES) return ECHO_ERR_UNSUPPORTED;
  if (((uintptr_t)d->q % 16) || ((uintptr_t)d->out % 16) || ((uintptr_t)d->gate % 16) || (d->heads * HD) % 8)
    return ECHO_ERR_UNSUPPORTED;
  Maps M;
  AttnParams A;
  memset(&A, 0, sizeof(A));
  memset(&M, 0, sizeof(M));
  {
    cuuint64_t dims[4] = {(cuuint64_t)HD, (cuuint64_t)d->heads, (cuuint64_t)d->seq, (cuuint64_t)d->rows};
    cuuint64_t str[3] = {(cuuint64_t)HD * 2, (cuuint64_t)d->q_row_stride * 2, (cuuint64_t)d->q_batch_stride * 2};
    cuuint32_t box[4] = {64, 1, QT, 1};
    int rc = echo_make_tmap_bf16(&M.q, d->q, 4, dims, str, box);
    if (rc) return rc;
  }
  {
    const long long width = (long long)d->heads * HD;
    cuuint64_t dims[4] = {(cuuint64_t)HD, (cuuint64_t)d->heads, (cuuint64_t)d->seq, (cuuint64_t)d->rows};
    cuuint64_t str[3] = {(cuuint64_t)HD * 2, (cuuint64_t)width * 2, (cuuint64_t)d->seq * width * 2};
    cuuint32_t box[4] = {64, 1, QT, 1};
    int rc = echo_make_tmap_bf16(&M.out, d->out, 4, dims, str, box);
    if (rc) return rc;
    if (d->gate) {
      rc = echo_make_tmap_bf16(&M.gate, d->gate, 4, dims, str, box);
      if (rc) return rc;
    } else {
      M.gate = M.out;
    }
  }
#ifdef ECHO_DEBUG
  A.trace = g_attn_trace;
#endif
  A.static_first = (d->n_seg > 0 && d->seg[0].len > 0 && !d->seg[0].mask && !d->seg[0].extent) ? 1 : 0;
  A.rows = d->rows; A.seq = d->seq; A.heads = d->heads; A.n_seg = d->n_seg;
  A.softmax_scale = d->softmax_scale;
  A.gate = (const bf16*)d->gate; A.out = (bf16*)d->out;
  for (int i = 0; i < d->n_seg; ++i) {
    const echo_attn_segment& s = d->seg[i];
    A.seg[i].len = s.len; A.seg[i].kv_batch_mod = s.kv_batch_mod;
    A.seg[i].causal = s.causal ? 1 : 0; A.seg[i].q_offset = s.causal ? s.q_offset : 0;
    A.seg[i].mask = s.mask; A.seg[i].mask_stride = s.mask_stride;
    A.seg[i].extent = s.extent; A.seg[i].kv_scale = s.kv_scale;
    if (s.len <= 0) { M.k[i] = M.q; M.v[i] = M.q; continue; }
    if (((uintptr_t)s.k % 16) || ((uintptr_t)s.v % 16)) return ECHO_ERR_UNSUPPORTED;
    const int kv_rows = s.kv_batch_mod > 0 ? s.kv_batch_mod : d->rows;
    cuuint64_t dims[4] = {(cuuint64_t)HD, (cuuint64_t)d->heads, (cuuint64_t)s.len, (cuuint64_t)kv_rows};
    cuuint64_t str[3] = {(cuuint64_t)HD * 2, (cuuint64_t)s.row_stride * 2,
                         (cuuint64_t)(kv_rows > 1 ? s.batch_stride : (long long)s.len * s.row_stride) * 2};
    cuuint32_t box[4] = {64, 1, KT, 1};
    int rc = echo_make_tmap_bf16(&M.k[i], s.k, 4, dims, str, box);
    if (rc) return rc;
    rc = echo_make_tmap_bf16(&M.v[i], s.v, 4, dims, str, box);
    if (rc) return rc;
  }
  {   // the attribute is per device: a process that drives several GPUs sets it once on each
    static bool attr_set[64] = {};
    int dev = 0;
    ECHO_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
      ECHO_CUDA(cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_ATTN));
      if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
  }
  // split-KV when the (row, head, query tile) groups alone leave SMs idle (two CTAs fit one SM): the un-guided 640-token
  // step has 80 groups, a 160-latent streaming block 32.  The caller lends the exchange buffer.
  const int qtiles = ceil_div(d->seq, QT);
  const long long groups = (long long)qtiles * d->heads * d->rows;
  int splits = 1;
  if (d->kv_splits != 1 && d->workspace && groups > 0 && 2 * groups <= WS_COUNTERS) {
    static const int forced = [] { const char* e = getenv("ECHO_ATTN_SPLITS"); return e && e[0] ? atoi(e) : 0; }();
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int slots = 2 * sms;
    int want = d->kv_splits > 1 ? d->kv_splits : (forced > 0 ? forced : (int)(slots / groups));
    if (d->kv_splits <= 1 && forced <= 0 && want < MAX_SPLITS) want = 1;   // measured: fewer than four parts do not pay
                                                                            // for the exchange (the 640-token step: 80 groups)
    if (want > MAX_SPLITS) want = MAX_SPLITS;
    if (want > total_tiles / 2) want = (int)(total_tiles / 2);          // at least two key tiles per part
    const long long room = (d->workspace_bytes - (long long)WS_COUNTERS * 4) / ((long long)PART_FLOATS * 4);
    if (want > 1 && room >= groups * want) splits = want;
  }
  A.splits = splits;
  A.ws_count = reinterpret_cast<int*>(d->workspace);
  A.ws_parts = d->workspace ? reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(d->workspace) + (long long)WS_COUNTERS * 4) : nullptr;
  dim3 grid(qtiles * splits, d->heads, d->rows);
  ECHO_CUDA(echo_launch(attention_tc_kernel, grid, dim3(192), SMEM_ATTN, st, M, A));
  ECHO_COUNT_LAUNCH();
  ECHO_LAUNCH_CHECK();
  return ECHO_OK;
}
