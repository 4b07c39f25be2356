// bf16 GEMM / implicit causal conv for sm_100a -- host side of the one GEMM kernel of this library, the CTA-pair
// tcgen05 kernel of gemm2.cu: argument checking, descriptor -> kernel parameters, tile-width choice, tensor maps.
#include "common.cuh"
#include "ptx_sm100.cuh"
#include "gemm_epilogue.cuh"

constexpr int BM = 128;
constexpr int BK = 64;

// -------------------------------------------------------------------------------------------
// host side: tensor maps through the driver entry point (no link-time libcuda dependency)
// -------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  }
  return fn;
}

static int make_tmap(CUtensorMap* m, CUtensorMapDataType dt, const void* base, int rank, const cuuint64_t* dims,
                     const cuuint64_t* strides_b, const cuuint32_t* box, CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B);
// bf16 output tiles of the row epilogue (box rows of 32 elements = 64 B, staged unswizzled)
int echo_make_tmap_bf16_linear(CUtensorMap* m, const void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides_b,
                               const cuuint32_t* box) {
  return make_tmap(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, base, rank, dims, strides_b, box, CU_TENSOR_MAP_SWIZZLE_NONE);
}
int echo_make_tmap_bf16(CUtensorMap* m, const void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides_b,
                        const cuuint32_t* box) {
  return make_tmap(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, base, rank, dims, strides_b, box);
}
// fp32 output tiles of the TMA epilogue (box rows of 32 floats = 128 B, SWIZZLE_128B)
int echo_make_tmap_f32(CUtensorMap* m, const void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides_b,
                       const cuuint32_t* box) {
  return make_tmap(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, base, rank, dims, strides_b, box);
}
static int make_tmap(CUtensorMap* m, CUtensorMapDataType dt, const void* base, int rank, const cuuint64_t* dims,
                     const cuuint64_t* strides_b, const cuuint32_t* box, CUtensorMapSwizzle swz) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) { echo_set_error("cuTensorMapEncodeTiled entry point unavailable"); return ECHO_ERR_CUDA; }
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = enc(m, dt, rank, const_cast<void*>(base), dims, strides_b, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { echo_set_error("cuTensorMapEncodeTiled failed (%d)", (int)r); return ECHO_ERR_CUDA; }
  return ECHO_OK;
}

int echo_launch_gemm_pair(int epi, int bn, int splits, long long split_stride, void* workspace, long long workspace_bytes,
                          const GemmShape& G, const EpiParams& P, int batch, cudaStream_t st, int sched_flags);  // gemm2.cu; bit 0 stream-K allowed, bit 1 no narrow last round, bit 2 no row epilogue

static int fill_params(const echo_gemm_desc* d, GemmShape& G, EpiParams& P) {
  ECHO_CHECK_ARG(d && d->a && d->b, "gemm: null operand");
  ECHO_CHECK_ARG(d->rows > 0 && d->batch > 0 && d->n > 0 && d->k_tap > 0, "gemm: bad shape");
  ECHO_CHECK_ARG(d->taps >= 1 && d->taps <= 8, "gemm: taps must be 1..8");
  ECHO_CHECK_ARG(d->k_tap % 8 == 0 && d->lda % 8 == 0 && d->ldb % 8 == 0, "gemm: k/lda/ldb must be multiples of 8");
  ECHO_CHECK_ARG(d->n % 8 == 0, "gemm: n must be a multiple of 8");
  ECHO_CHECK_ARG(((uintptr_t)d->a % 16) == 0 && ((uintptr_t)d->b % 16) == 0, "gemm: operands must be 16-byte aligned");
  ECHO_CHECK_ARG(d->ldb >= (long long)d->taps * d->k_tap, "gemm: ldb < taps*k_tap");
  memset(&G, 0, sizeof(G));
  memset(&P, 0, sizeof(P));
  G.k_blocks_per_tap = ceil_div(d->k_tap, BK);
  G.taps = d->taps; G.k_tap = d->k_tap;
  for (int i = 0; i < 8; ++i) G.tap_shift[i] = d->tap_shift[i];
  G.rows = d->rows; G.n = d->n; G.lda = d->lda; G.a_batch_stride = d->a_batch_stride; G.ldb = d->ldb;
  G.a = (const bf16*)d->a; G.b = (const bf16*)d->b;
  G.b_static = d->b_static == 1;
  ECHO_CHECK_ARG(d->a_ctx >= 0 && d->a_ctx < (1 << 20), "gemm: bad a_ctx");
  if (d->a_ctx > 0) {   // left context: the operand view starts a_ctx rows earlier, every tap moves with it
    G.a_ctx = d->a_ctx;
    G.a -= (long long)d->a_ctx * d->lda;
    for (int i = 0; i < 8; ++i) G.tap_shift[i] += d->a_ctx;
    ECHO_CHECK_ARG(((uintptr_t)G.a % 16) == 0, "gemm: context rows must keep the operand 16-byte aligned");
    ECHO_CHECK_ARG(d->batch == 1 || d->a_batch_stride >= (long long)(d->rows + d->a_ctx) * d->lda, "gemm: a_batch_stride must span the context rows");
  }
  ECHO_CHECK_ARG((long long)d->rows * d->batch < (1LL << 31), "gemm: rows * batch must fit 31 bits");
  P.kind = d->epilogue; P.rows = d->rows; P.n = d->n; P.act = d->act;
  P.bias = d->bias; P.gate = d->gate; P.gate_ld = d->gate_ld; P.rows_per_gate = d->rows_per_gate > 0 ? d->rows_per_gate : 1;
  P.resid = d->resid; P.ld_resid = d->ld_resid; P.out_f32 = d->out_f32; P.ld_out_f32 = d->ld_out_f32;
  P.red_resid = (d->resid && d->resid == d->out_f32 && d->ld_resid == d->ld_out_f32 && !d->out_bf16 &&
                 !d->out2_bf16 && d->k_splits <= 1) ? 1 : 0;
  P.out_bf16 = (bf16*)d->out_bf16; P.ld_out_bf16 = d->ld_out_bf16;
  P.out2_bf16 = (bf16*)d->out2_bf16; P.ld_out2_bf16 = d->ld_out2_bf16; P.snake_alpha = d->snake_alpha;
  P.n_sections = d->n_sections; P.section_width = d->section_width; P.head_dim = d->head_dim;
  P.rope_heads = d->rope_heads; P.pos_offset = d->pos_offset; P.pos_stride = d->pos_stride;
  P.rows_per_seq = d->rows_per_seq > 0 ? d->rows_per_seq : 1; P.eps = d->eps;
  P.rope_cos = d->rope_cos; P.rope_sin = d->rope_sin; P.ld_sec_out = d->ld_sec_out;
  for (int i = 0; i < 4; ++i) {
    P.sec[i].type = d->sections[i].type; P.sec[i].out = (bf16*)d->sections[i].out;
    P.sec[i].norm_w = d->sections[i].norm_weight;
  }
  {   // vectorised epilogue: 16-byte loads of the per-column vectors
    auto al16 = [](const void* p) { return ((uintptr_t)p % 16) == 0; };
    ECHO_CHECK_ARG(al16(d->bias) && al16(d->gate) && al16(d->snake_alpha) && al16(d->resid) && al16(d->out_f32) &&
                   d->gate_ld % 4 == 0, "gemm: bias/gate/alpha/resid/out_f32 must be 16-byte aligned (gate_ld %% 4)");
    ECHO_CHECK_ARG(((uintptr_t)d->out_bf16 % 8) == 0 && ((uintptr_t)d->out2_bf16 % 8) == 0 &&
                   ((uintptr_t)d->rope_cos % 8) == 0 && ((uintptr_t)d->rope_sin % 8) == 0, "gemm: bf16 outputs / rope tables must be 8-byte aligned");
    for (int i = 0; i < 4; ++i)
      ECHO_CHECK_ARG(al16(d->sections[i].norm_weight) && ((uintptr_t)d->sections[i].out % 8) == 0, "gemm: section pointers misaligned");
  }
  if (d->epilogue == ECHO_EPI_GENERIC) {
    ECHO_CHECK_ARG(d->out_f32 || d->out_bf16 || d->out2_bf16, "gemm: no output");
    ECHO_CHECK_ARG(!d->out2_bf16 || d->snake_alpha, "gemm: out2 needs snake_alpha");
    ECHO_CHECK_ARG((!d->out_f32 || d->ld_out_f32 % 4 == 0) && (!d->resid || d->ld_resid % 4 == 0) &&
                   (!d->out_bf16 || d->ld_out_bf16 % 8 == 0) && (!d->out2_bf16 || d->ld_out2_bf16 % 8 == 0),
                   "gemm: output leading dims must keep 16-byte alignment");
  } else if (d->epilogue == ECHO_EPI_SECTIONS) {
    ECHO_CHECK_ARG(d->n_sections >= 1 && d->n_sections <= 4 && d->n == d->n_sections * d->section_width,
                   "gemm: sections do not tile n");
    ECHO_CHECK_ARG(d->head_dim == 64 || d->head_dim == 128, "gemm: head_dim must be 64 or 128");
    ECHO_CHECK_ARG(d->section_width % d->head_dim == 0 && d->ld_sec_out % 8 == 0, "gemm: bad section layout");
    for (int i = 0; i < d->n_sections; ++i) {
      ECHO_CHECK_ARG(d->sections[i].out, "gemm: section output missing");
      ECHO_CHECK_ARG(d->sections[i].type != ECHO_SEC_NORM_ROPE || d->sections[i].norm_weight, "gemm: norm weight missing");
      ECHO_CHECK_ARG((d->sections[i].type != ECHO_SEC_NORM_ROPE && d->sections[i].type != ECHO_SEC_ROPE) ||
                     d->rope_heads == 0 || (d->rope_cos && d->rope_sin), "gemm: rope table missing");
    }
  } else if (d->epilogue == ECHO_EPI_SWIGLU) {
    ECHO_CHECK_ARG(d->out_bf16 && d->n % 128 == 0 && d->ld_out_bf16 % 8 == 0, "gemm: swiglu needs out_bf16, n%%128==0");
  } else {
    ECHO_CHECK_ARG(false, "gemm: unknown epilogue %d", d->epilogue);
  }
  return ECHO_OK;
}

// Tile width for the CTA-pair kernel: fewest (rounds x tile cost) over the 74 pairs; ties go to the wider tile.
static int pick_block_n_pair(const echo_gemm_desc* d) {
  if (d->block_n) return d->block_n;
  if (d->k_splits > 1) return d->n >= 256 ? 256 : ((d->n + 31) / 32) * 32;   // split-K callers size splits for 256-wide tiles
  static int npairs = 0;
  if (!npairs) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    npairs = sms > 1 ? sms / 2 : 74;
  }
  const int cand_generic[] = {256, 192, 128, 96, 64};
  const int cand_fused[] = {256, 128, 64};
  const int* cand = d->epilogue == ECHO_EPI_GENERIC ? cand_generic : cand_fused;
  const int ncand = d->epilogue == ECHO_EPI_GENERIC ? 5 : 3;
  const long long tm = (long long)ceil_div(d->rows, 2 * BM) * d->batch;
  int best = 0;
  long long best_cost = 0;
  for (int i = 0; i < ncand; ++i) {
    const int bn = cand[i];
    if (d->epilogue == ECHO_EPI_SECTIONS && (bn % d->head_dim != 0 || d->section_width % bn != 0)) continue;
    if (d->epilogue == ECHO_EPI_SWIGLU && bn % 128 != 0) continue;
    const long long tiles = tm * ceil_div(d->n, bn);
    const long long rounds = (tiles + npairs - 1) / npairs;
    // measured (tools/gemm_trace.py): a k-block costs ~600 clocks whether the tile is 128 or 256 columns wide (operand
    // feed, not MMA issue, is the floor), so narrow tiles only pay off when they remove a whole round or shorten the
    // exposed epilogue of a single-round launch.
    const long long cost = rounds * (bn > 240 ? bn : 240) + bn / 2;
    if (!best || cost < best_cost) { best = bn; best_cost = cost; }
  }
  return best;
}

extern "C" int echo_gemm_bf16(const echo_gemm_desc* d, echo_stream_t stream) {
  GemmShape G;
  EpiParams P;
  int rc = fill_params(d, G, P);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  ECHO_CHECK_ARG(d->impl == 0 || (d->impl >= 2 && d->impl <= 5),
                 "gemm: impl %d: only the CTA-pair tcgen05 kernel exists (0 / 2; 3 = stream-K schedule allowed; 4 = plain schedule without the narrow last round; 5 = transposing GENERIC epilogue instead of the row / TMA one)", d->impl);
  const int bn = pick_block_n_pair(d);
  ECHO_CHECK_ARG(bn > 0, "pair gemm: no tile width fits epilogue %d (n=%d)", d->epilogue, d->n);
  if (d->epilogue == ECHO_EPI_SECTIONS)
    ECHO_CHECK_ARG(d->section_width % bn == 0 && bn % d->head_dim == 0, "gemm: block_n %d does not tile sections", bn);
  if (d->epilogue == ECHO_EPI_SWIGLU) ECHO_CHECK_ARG(bn % 128 == 0, "gemm: swiglu needs block_n %% 128 == 0");
  if (d->k_splits > 1)
    ECHO_CHECK_ARG(d->epilogue == ECHO_EPI_GENERIC && d->act == ECHO_ACT_NONE && !d->resid && d->out_f32 && !d->out_bf16 &&
                   !d->out2_bf16 && d->split_stride % 4 == 0, "gemm: k_splits needs GENERIC, no act/resid, fp32 output only");
  return echo_launch_gemm_pair(d->epilogue, bn, d->k_splits, d->split_stride, d->workspace, d->workspace_bytes, G, P,
                               d->batch, st, (d->impl == 3 ? 1 : 0) | (d->impl == 4 ? 2 : 0) | (d->impl == 5 ? 4 : 0));
}
