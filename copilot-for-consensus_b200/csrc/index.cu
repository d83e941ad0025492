// index.cu -- host side of the vector index behind the C ABI (include/b200_retrieval.h):
// storage, ingest, search orchestration, compaction.  Device code is in search_kernels.cuh.
//
// Replaces the storage + arithmetic of InMemoryVectorStore
// (adapters/copilot_vectorstore/copilot_vectorstore/inmemory.py:14-178); ids, metadata and the
// Python exception contract live in the wrapper (vectorstore/store.py).
//
// Data layout in HBM (one index = one GPU's shard):
//   search  bf16 [capacity, dpad]  unit-normalised rows, dpad = dim rounded up to 64; the operand
//                                  the tcgen05 tiles stream through TMA (768 B/row at dim 384)
//   master  fp32|bf16 [capacity, dim]  the caller's rows, untouched: SearchResult.vector and the
//                                  exact fp32 rescoring of the KP survivors per query
//   cand    u64 [nq, nsplit, KP]   per-split candidate keys (scratch between K8 and K9)
#include <algorithm>
#include <atomic>
#include <vector>

#include "host_util.h"
#include "search_kernels.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------
// error + driver plumbing (shared with encoder.cu through host_util.h)
// ------------------------------------------------------------------------------------------
static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* get_error() { return g_err; }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = []() -> EncodeTiledFn {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

int make_tma_2d_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows,
                     uint32_t box_cols) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) B200_FAIL(B200_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the CUDA driver");
  if (box_cols * 2 != 128 || box_rows > 256) B200_FAIL(B200_ERR_INVALID, "bad TMA box %u x %u", box_rows, box_cols);
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {cols * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    B200_FAIL(B200_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (rows %llu cols %llu)", (int)r,
              (unsigned long long)rows, (unsigned long long)cols);
  return B200_OK;
}

int make_tma_2d_bf16_sw64(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows,
                          uint32_t box_cols) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) B200_FAIL(B200_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the CUDA driver");
  if (box_cols * 2 != 64 || box_rows > 256) B200_FAIL(B200_ERR_INVALID, "bad TMA box %u x %u", box_rows, box_cols);
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {cols * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) B200_FAIL(B200_ERR_CUDA, "cuTensorMapEncodeTiled (64B swizzle) failed with CUresult %d", (int)r);
  return B200_OK;
}

int make_tma_2d_bf16_plain(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows,
                           uint32_t box_cols) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) B200_FAIL(B200_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the CUDA driver");
  if ((box_cols * 2) % 16 != 0 || box_cols > 256 || box_rows > 256) B200_FAIL(B200_ERR_INVALID, "bad TMA box %u x %u", box_rows, box_cols);
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {cols * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) B200_FAIL(B200_ERR_CUDA, "cuTensorMapEncodeTiled (no swizzle) failed with CUresult %d", (int)r);
  return B200_OK;
}

int check_device_is_sm100(int device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    cudaGetLastError();
    B200_FAIL(B200_ERR_CUDA, "no CUDA device is visible; this library has no CPU fallback");
  }
  if (device < 0 || device >= n) B200_FAIL(B200_ERR_INVALID, "device %d out of range (0..%d)", device, n - 1);
  int major = 0;
  B200_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  if (major != 10)
    B200_FAIL(B200_ERR_UNSUPPORTED, "device %d has compute capability %d.x; the kernels are sm_100a only", device,
              major);
  return B200_OK;
}

// ---- kernel-variant switches (b200_set_option) ---------------------------------------------------
struct OptDef { const char* name; int def, lo, hi; };
static const OptDef kOptDefs[OPT_COUNT] = {
    {"pdl", 0, 0, 1},           {"gemm_pair", 1, 0, 1},      {"gemm_ast", 1, 0, 1},   {"attention_tc", 1, 0, 1},
    {"att_exact_max", 0, 0, 1}, {"att_turns", 1, 0, 1},      {"att_poly", 1, 0, 4},   {"fuse_ln", 1, 0, 1},
    {"fuse_ffn", 1, 0, 1},      {"serpentine", 1, 0, 1},     {"sim_pair", 1, 0, 1},   {"sim_bound", 3, 0, 3},
    {"sim_refresh_mask", 0, 0, 1023}, {"search_certify", 1, 0, 1}, {"sim_pair_kp", 0, 0, 20}, {"sim_nsplit", 0, 0, 4096},
    {"search_screen", 1, 0, 1}, {"att_norms_fused", 1, 0, 1}, {"defer_ln", 1, 0, 1}, {"sim_drain", 0, 0, 2},
};
static std::atomic<int> g_opts[OPT_COUNT];
static std::once_flag g_opts_once;
static void opts_init() {
  std::call_once(g_opts_once, [] { for (int i = 0; i < OPT_COUNT; ++i) g_opts[i].store(kOptDefs[i].def, std::memory_order_relaxed); });
}
int opt(Option o) {
  opts_init();
  return g_opts[o].load(std::memory_order_relaxed);
}

// ---- optional per-kernel timing ----------------------------------------------------------------
struct ProfRec {
  const char* name;
  double flops, bytes;
  cudaEvent_t a, b;
};
static std::mutex g_prof_mu;
static std::vector<ProfRec> g_prof;
static bool g_prof_on = false;

bool prof_enabled() { return g_prof_on; }

ProfScope::ProfScope(const char* name, double flops, double bytes, cudaStream_t s) : st(s) {
  if (!g_prof_on) return;
  ProfRec r{name, flops, bytes, nullptr, nullptr};
  if (cudaEventCreate(&r.a) != cudaSuccess || cudaEventCreate(&r.b) != cudaSuccess) return;
  cudaEventRecord(r.a, st);
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof.push_back(r);
  slot = (int)g_prof.size() - 1;
}
ProfScope::~ProfScope() {
  if (slot < 0) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (slot < (int)g_prof.size()) cudaEventRecord(g_prof[slot].b, st);
}

}  // namespace b200

using namespace b200;

struct b200_index {
  int dim = 0, dpad = 0, device = 0, master_dtype = 0, num_sms = 148;
  int64_t count = 0, capacity = 0;
  DevBuf search, master, stage, q16, q_f32, qmeta, cand, gbkt, out_ids, out_scores, tmp;
  // certificate state: persist[0] = rvmax (float bits), persist[2..3] / [4..5] = the two 64-bit counters
  // of b200_index_certificate_stats; cert = [fail_count][pad to 64 B][ticket nq][fail_idx nq] per search
  DevBuf persist, cert, exact_part;
  int exact_per_sm[3] = {0, 0, 0};   // resident CTAs per SM of exact_topk_kernel<0,0 / 1,0 / 0,1> at this index's dimension
  StreamOrder order;
  std::mutex mu;
  size_t master_elt() const { return master_dtype == B200_INDEX_MASTER_BF16 ? 2 : 4; }
  uint32_t* rvmax() const { return persist.as<uint32_t>(); }
  unsigned long long* stats() const { return reinterpret_cast<unsigned long long*>(persist.as<uint8_t>() + 8); }
};

static constexpr int64_t kMaxRows = (int64_t)1 << 31;  // row ids are uint32 on the device, 2^32-1 is "no row"
static constexpr int kMaxQueryChunk = 8192;
// Candidates kept per query and CTA for top_k <= 10.  The certificate fails when the k-th exact score and the KP-th
// 16-bit score are closer than the rounding bound eps (~7e-4): on random 384-d rows that is a sum of KP - k order-
// statistic spacings, so its probability falls like eps^(KP - k): measured 2.4e-4 per query at KP = 16, none in
// 10^5 queries at KP = 20.  A single query tile (<= 128 queries, the HBM-bound regime) has the issue slots to spare
// for 20; the CTA-pair kernel (tensor-bound; the larger sets cost it one of its seven pipeline stages) slows down with
// 20 by 0.54 ms at 1.25 M rows (19 %), 0.2 ms at 5 - 10 M (1 - 2 %), 0.4 ms at 25 M, 3.2 ms at 50 M (4096 queries),
// while 16 sends 1 - 2 queries of every 4096 to the screened exhaustive pass (0.03 ms + 0.17 - 0.23 ms per million rows
// each: 0.3 ms at 1.25 M, 1.7 ms at 10 M, 11.3 ms at 50 M) -- measured in bench.py's 10 M-row leg: 2 re-runs per batch,
// 27.9 ms against 24.2.  The two costs cross near 2 M rows per shard: 16 below kPairKp20FromRows, 20 from there on.
static constexpr int kKpSmall = 20;
static constexpr int kKpPair = 16;
static constexpr int64_t kPairKp20FromRows = 2500000;
static constexpr int kMaxCandPerQuery = 8192;
// shards below this many rows run the CTA-pair similarity kernel with two drain groups (see sim_topk_kernel, DR)
static constexpr int64_t kDualDrainBelowRows = 2000000;

static int index_reserve_locked(b200_index* ix, int64_t cap, cudaStream_t st) {
  if (cap <= ix->capacity) return B200_OK;
  if (cap > kMaxRows) B200_FAIL(B200_ERR_UNSUPPORTED, "index capacity %lld exceeds %lld rows per GPU", (long long)cap, (long long)kMaxRows);
  int64_t want = std::max<int64_t>(cap, std::min<int64_t>(kMaxRows, ix->capacity + ix->capacity / 2));
  want = (want + 255) / 256 * 256;
  B200_TRY(ix->search.grow((size_t)want * ix->dpad * 2, (size_t)ix->count * ix->dpad * 2, st));
  B200_TRY(ix->master.grow((size_t)want * ix->dim * ix->master_elt(), (size_t)ix->count * ix->dim * ix->master_elt(), st));
  ix->capacity = want;
  return B200_OK;
}

// CG = 2 runs CTA pairs (two query tiles per cluster); nqt is then the even number of query tiles.
// DR = 2: two drain groups per CTA (see sim_topk_kernel); the launch then hands K9 nsplit * DR candidate lists per query,
// which is what `used->nsplit` reports.
template <int BLOCK_N, int KP, int CG, int DR = 1>
static int launch_sim(b200_index* ix, int nq, int nqt, const CUtensorMap& map_q, cudaStream_t st, SimParams* used) {
  const int kc = ix->dpad / SIM_BLOCK_K;
  const size_t a_bytes = (size_t)kc * SIM_A_CHUNK_BYTES;
  const size_t list_bytes = (size_t)DR * SIM_BLOCK_M * (KP + 1) * 8;
  const size_t stage_bytes = (size_t)(BLOCK_N / CG) * SIM_BLOCK_K * 2;
  size_t fixed = 1024 + a_bytes + list_bytes + 256;
  size_t stage_total = stage_bytes;
  const size_t max_smem = 232448;  // 227 KB opt-in limit per CTA
  const int units = nqt / CG;              // CTA groups per corpus slice
  const int slots = ix->num_sms / CG;      // CTA groups resident at once
  // dim <= 384: the 128-query tile stays in shared memory.  Larger dimensions: its 64-wide chunks travel through
  // the ring beside the corpus chunks (they come from L2: the tile is 128 x dpad bf16) -- 50 % more L2 -> SM
  // traffic per corpus tile, the same HBM traffic.
  const int a_stream = fixed + 4 * stage_bytes > max_smem ? 1 : 0;
  if (a_stream) {
    fixed = 1024 + list_bytes + 256;
    stage_total = stage_bytes + SIM_A_CHUNK_BYTES;
  }
  int nstage = (int)std::min<size_t>(8, (max_smem - fixed) / stage_total);
  const size_t smem = fixed + (size_t)nstage * stage_total;

  const uint32_t n_rows = (uint32_t)ix->count;
  const uint32_t total_tiles = (n_rows + BLOCK_N - 1) / BLOCK_N;
  // Split the corpus so that (query tiles x splits) fills whole waves of SMs.  Cost model: every
  // CTA walks tiles_per_split tiles plus a fixed prologue worth ~2 tiles (query-tile load, TMEM
  // allocation, list write-out); waves run back to back.
  const int max_split = std::max(1, std::min<int>((int)std::min<uint32_t>(total_tiles, 65535u), kMaxCandPerQuery / (KP * DR)));
  int best = 1;
  double best_cost = 1e300;
  for (int w = 1; w <= 12; ++w) {
    int ns = std::min(max_split, std::max(1, (slots * w) / units));
    const uint32_t t = (total_tiles + ns - 1) / ns;
    ns = (int)((total_tiles + t - 1) / t);
    if (w > 1 && t < 8) break;
    const long items = (long)ns * units;
    const long waves = (items + slots - 1) / slots;
    const double cost = (double)waves * ((double)t + 2.0);
    if (cost < best_cost * 0.995) { best_cost = cost; best = ns; }
  }
  int nsplit = best;
  if (opt(OPT_SIM_NSPLIT) > 0) nsplit = std::min(max_split, opt(OPT_SIM_NSPLIT));     // A/B timing
  const uint32_t tps = (total_tiles + nsplit - 1) / nsplit;
  nsplit = (int)((total_tiles + tps - 1) / tps);

  const size_t cand_n = (size_t)nqt * SIM_BLOCK_M * nsplit * KP * DR;
  B200_TRY(ix->cand.grow(cand_n * 8, 0, st));
  // [nq_pad][KP] bucket bests followed by [nq_pad] single-CTA bounds; both start at "none"
  const size_t bkt_words = (size_t)nqt * SIM_BLOCK_M * (KP + 1);
  B200_TRY(ix->gbkt.grow(bkt_words * 4, 0, st));
  B200_CUDA(cudaMemsetAsync(ix->gbkt.p, 0, bkt_words * 4, st));

  CUtensorMap map_c;
  B200_TRY(make_tma_2d_bf16(&map_c, ix->search.p, (uint64_t)ix->count, (uint64_t)ix->dpad, BLOCK_N / CG, SIM_BLOCK_K));

  SimParams p;
  p.nq = nq; p.kc = kc; p.nstage = nstage; p.n_rows = n_rows; p.tiles_per_split = tps; p.nsplit = nsplit;
  p.a_stream = a_stream;
  p.cand = ix->cand.as<uint64_t>(); p.gbkt = ix->gbkt.as<uint32_t>(); p.use_bound = nsplit >= KP;
  p.gthr = ix->gbkt.as<uint32_t>() + (size_t)nqt * SIM_BLOCK_M * KP;
  p.refresh_mask = (uint32_t)opt(OPT_SIM_REFRESH_MASK);
  {
    // bit 0: the bucket bound (needs one bucket per kept candidate); bit 1: the single-CTA stand-in bound, which
    // only earns its atomics while buckets are still unreported, i.e. when the launch takes more than one wave
    const int m = opt(OPT_SIM_BOUND);
    p.bkt_per_cta = nsplit >= KP ? 1 : (KP + nsplit - 1) / nsplit;
    p.use_bound = (m & 1) && (int)tps >= p.bkt_per_cta;
    p.use_gthr = ((m >> 1) & 1) && ((long)nsplit * units > slots || !p.use_bound);
  }
  auto kern = sim_topk_kernel<BLOCK_N, KP, CG, DR>;
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  {
    // algorithmic work of one pass: every corpus row read once (768 B at dim 384) and 2*dpad flops
    // per (query, row) pair
    B200_PROF("sim_topk", 2.0 * nq * (double)n_rows * ix->dpad, (double)n_rows * ix->dpad * 2.0, st);
    if (CG == 2) {
      cudaLaunchConfig_t cfg{};
      cfg.gridDim = dim3(nqt, nsplit);          // consecutive blockIdx.x = the two CTAs of a pair
      cfg.blockDim = dim3(sim_threads(DR));
      cfg.dynamicSmemBytes = smem;
      cfg.stream = st;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr; cfg.numAttrs = 1;
      B200_CUDA(cudaLaunchKernelEx(&cfg, kern, map_q, map_c, p));
    } else {
      kern<<<dim3(nqt, nsplit), sim_threads(DR), smem, st>>>(map_q, map_c, p);
    }
  }
  B200_CUDA(cudaGetLastError());
  *used = p;
  used->nsplit = nsplit * DR;      // candidate lists per query, as K9 sees them
  return B200_OK;
}

// bound on the evaluation error of the two scores the certificate compares: the tensor cores' fp32
// accumulation of d exact products (<= 2^-23 of the running sum per addition even when it truncates, and the sum
// of the magnitudes is <= 1 in cosine units) and the canonical fp32 routine
static float certificate_slop(int d) { return (float)d * 1.1920929e-7f + (float)(d / 32 + 16) * 1.1920929e-7f + 1e-6f; }

// exact_topk_kernel over the fail list of the K9 launch just issued (or, with fail == nullptr, over all nq queries)
template <int METRIC_L2, int SCREEN = 0>
static int launch_exact(b200_index* ix, const uint32_t* fail_count, const int32_t* fail_idx, uint32_t* ticket, int nq,
                        const float* q_f32, int k, int64_t id_base, int64_t* ids_dev, float* sc_dev, cudaStream_t st,
                        uint64_t* part = nullptr, int grid = 0, HitSink packed_dev = HitSink{},
                        uint32_t* fail2_count = nullptr, int32_t* fail2_idx = nullptr) {
  // shared memory: room to stage one 64-key list per CTA (one CTA per SM: the kernel is register-bound to that anyway)
  const int stage_keys = std::max(1024, ix->num_sms * EXACT_KEEP);
  const size_t smem = (size_t)stage_keys * 8 + (size_t)ix->dim * 4;
  auto kern = exact_topk_kernel<METRIC_L2, SCREEN>;
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // The kernel's CTAs wait for one another (a query's list buffer is reused two queries later), so every CTA of the
  // grid must be able to be resident at once: never launch more than the occupancy calculator allows.
  int& per_sm = ix->exact_per_sm[METRIC_L2 + 2 * SCREEN];
  if (per_sm == 0) B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, EXACT_THREADS, smem));
  if (per_sm < 1) B200_FAIL(B200_ERR_UNSUPPORTED, "the exhaustive kernel does not fit on an SM (dimension %d)", ix->dim);
  if (!part) {   // the index's own scratch (grown on first use); the no-allocation entry point passes its own
    grid = ix->num_sms;
    B200_TRY(ix->exact_part.grow((size_t)2 * grid * EXACT_KEEP * 8, 0, st));
    part = ix->exact_part.as<uint64_t>();
  }
  grid = std::min(grid, per_sm * ix->num_sms);
  ExactParams e{};
  e.fail_count = fail_count; e.fail_idx = fail_idx; e.nq = nq; e.q_f32 = q_f32;
  e.master = ix->master.p; e.master_is_bf16 = ix->master_dtype == B200_INDEX_MASTER_BF16;
  e.n_rows = (uint32_t)ix->count; e.d = ix->dim; e.k = k; e.id_base = id_base;
  e.part = part; e.stage_keys = stage_keys; e.ticket = ticket; e.out_ids = ids_dev; e.out_scores = sc_dev; e.out_packed = packed_dev;
  e.fail2_count = fail2_count; e.fail2_idx = fail2_idx;
  // screen and canonical score are two fp32 evaluations of the same number: apart by at most a few d * 2^-24
  e.screen_margin = 1e-4f + (float)ix->dim * 4.8e-7f;
  {
    B200_PROF(METRIC_L2 ? "exact_l2" : (SCREEN ? "exact_topk" : "exact_full"), 0.0, 0.0, st);
    kern<<<grid, EXACT_THREADS, smem, st>>>(e);
  }
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

static HitSink sink_at(const HitSink& s, size_t off) {
  HitSink r = s;
  if (r.one) r.one += off;
  for (int i = 0; i < r.n_peers; ++i) r.peer[i] += off;
  return r;
}
static bool sink_on(const HitSink& s) { return s.one != nullptr || s.n_peers > 0; }

static int search_chunk(b200_index* ix, const void* q_dev, int q_dtype, int nq, int k, int kp, int64_t id_base,
                        int64_t* ids_dev, float* sc_dev, HitSink packed_dev, cudaStream_t st) {
  int nqt = (nq + SIM_BLOCK_M - 1) / SIM_BLOCK_M;
  // CTA pairs (two query tiles per cluster) for k <= 10 once there are at least two query tiles; an odd
  // tile count is padded with an empty tile.  Option sim_pair=0 forces single-CTA tiles (A/B timing).
  const bool pair = kp == kKpSmall && nqt >= 2 && opt(OPT_SIM_PAIR) != 0;
  // 16 candidates in the CTA-pair kernel unless the shard is so large that the exhaustive re-runs cost more than the
  // slower drain (see kPairKp20FromRows)
  if (pair && ix->count < kPairKp20FromRows) kp = kKpPair;
  if (pair && opt(OPT_SIM_PAIR_KP) == kKpPair) kp = kKpPair;      // A/B timing
  if (pair && opt(OPT_SIM_PAIR_KP) == kKpSmall) kp = kKpSmall;
  if (pair) nqt = (nqt + 1) & ~1;
  const int nq_pad = nqt * SIM_BLOCK_M;
  B200_TRY(ix->q16.grow((size_t)nq_pad * ix->dpad * 2, 0, st));
  B200_TRY(ix->q_f32.grow((size_t)nq * ix->dim * 4, 0, st));
  B200_TRY(ix->qmeta.grow((size_t)nq * 8, 0, st));
  const bool certify = opt(OPT_SEARCH_CERTIFY) != 0;
  // cert = [fail count, second fail count, pad to 64 B][tickets nq][tickets of the second pass nq][fail list nq][second fail list nq]
  const size_t cert_bytes = 64 + (size_t)nq * 16;
  if (certify) {
    B200_TRY(ix->cert.grow(cert_bytes, 0, st));
    B200_CUDA(cudaMemsetAsync(ix->cert.p, 0, 64 + (size_t)nq * 8, st));   // the counts and the tickets
  }
  const int wpb = 8;
  if (q_dtype == 0)
    prep_queries_kernel<float><<<(nq_pad + wpb - 1) / wpb, wpb * 32, 0, st>>>(
        (const float*)q_dev, nq, nq_pad, ix->dim, ix->dpad, ix->q16.as<__half>(), ix->q_f32.as<float>(), ix->qmeta.as<float2>());
  else
    prep_queries_kernel<__nv_bfloat16><<<(nq_pad + wpb - 1) / wpb, wpb * 32, 0, st>>>(
        (const __nv_bfloat16*)q_dev, nq, nq_pad, ix->dim, ix->dpad, ix->q16.as<__half>(), ix->q_f32.as<float>(), ix->qmeta.as<float2>());
  B200_CUDA(cudaGetLastError());

  CUtensorMap map_q;
  B200_TRY(make_tma_2d_bf16(&map_q, ix->q16.p, (uint64_t)nq_pad, (uint64_t)ix->dpad, SIM_BLOCK_M, SIM_BLOCK_K));
  SimParams sp{};
  // two drain groups per CTA on small shards, where the warm-up of the candidate sets is a large share of the launch
  // (option sim_drain: 1 / 2 force it; A/B timing)
  const bool dual = pair && (opt(OPT_SIM_DRAIN) == 2 || (opt(OPT_SIM_DRAIN) == 0 && ix->count < kDualDrainBelowRows));
  if (dual && kp == kKpPair) B200_TRY((launch_sim<256, kKpPair, 2, 2>(ix, nq, nqt, map_q, st, &sp)));
  else if (dual) B200_TRY((launch_sim<256, kKpSmall, 2, 2>(ix, nq, nqt, map_q, st, &sp)));
  else if (pair && kp == kKpPair) B200_TRY((launch_sim<256, kKpPair, 2>(ix, nq, nqt, map_q, st, &sp)));
  else if (pair) B200_TRY((launch_sim<256, kKpSmall, 2>(ix, nq, nqt, map_q, st, &sp)));
  else if (kp == kKpSmall) B200_TRY((launch_sim<256, kKpSmall, 1>(ix, nq, nqt, map_q, st, &sp)));
  else if (kp == 32) B200_TRY((launch_sim<128, 32, 1>(ix, nq, nqt, map_q, st, &sp)));
  else B200_TRY((launch_sim<128, 64, 1>(ix, nq, nqt, map_q, st, &sp)));
  const int nsplit = sp.nsplit;

  MergeParams m;
  m.gbkt = sp.gbkt; m.gthr = sp.gthr; m.use_bound = sp.use_bound; m.use_gthr = sp.use_gthr;
  m.nq = nq; m.nsplit = nsplit; m.kp = kp; m.k = k; m.d = ix->dim;
  int pow2 = 64;
  while (pow2 < nsplit * kp) pow2 <<= 1;
  m.ncand_pow2 = pow2;
  m.n_rows = (uint32_t)ix->count; m.id_base = id_base;
  m.cand = ix->cand.as<uint64_t>();
  m.q_f32 = ix->q_f32.as<float>(); m.master = ix->master.p;
  m.master_is_bf16 = ix->master_dtype == B200_INDEX_MASTER_BF16;
  m.out_ids = ids_dev; m.out_scores = sc_dev; m.out_packed = packed_dev;
  uint32_t* fail_count = certify ? ix->cert.as<uint32_t>() : nullptr;
  uint32_t* ticket = certify ? ix->cert.as<uint32_t>() + 16 : nullptr;
  uint32_t* ticket2 = certify ? ticket + nq : nullptr;
  int32_t* fail_idx = certify ? reinterpret_cast<int32_t*>(ticket + 2 * (size_t)nq) : nullptr;
  int32_t* fail2_idx = certify ? fail_idx + nq : nullptr;
  uint32_t* fail2_count = certify ? ix->cert.as<uint32_t>() + 1 : nullptr;
  m.qmeta = ix->qmeta.as<float2>(); m.rvmax = ix->rvmax(); m.slop = certificate_slop(ix->dpad);
  m.fail_count = fail_count; m.fail_idx = fail_idx; m.stats = ix->stats();
  const size_t msmem = (size_t)pow2 * 8 + (size_t)((ix->dim + 1) & ~1) * 4 + 64 * 8;
  B200_CUDA(cudaFuncSetAttribute(merge_rescore_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)msmem));
  {
    B200_PROF("merge_rescore", 0.0, (double)nq * nsplit * kp * 8.0 + (double)nq * kp * ix->dim * ix->master_elt(), st);
    merge_rescore_kernel<<<nq, 256, msmem, st>>>(m);
  }
  B200_CUDA(cudaGetLastError());
  // Queries the certificate turned down are redone against every row; the kernels return at once when there are none.
  // First the screened pass (one cheap pass over the rows, canonical scores for the best 64 only), then the full one
  // for what the screened pass could not settle (large clusters of near-duplicates).
  if (certify) {
    const bool screen = (ix->dim & 7) == 0 && opt(OPT_SEARCH_SCREEN) != 0;
    if (screen) {
      B200_TRY((launch_exact<0, 1>(ix, fail_count, fail_idx, ticket, nq, ix->q_f32.as<float>(), k, id_base, ids_dev, sc_dev, st,
                                   nullptr, 0, packed_dev, fail2_count, fail2_idx)));
      B200_TRY((launch_exact<0, 0>(ix, fail2_count, fail2_idx, ticket2, nq, ix->q_f32.as<float>(), k, id_base, ids_dev, sc_dev, st,
                                   nullptr, 0, packed_dev)));
    } else {
      B200_TRY((launch_exact<0, 0>(ix, fail_count, fail_idx, ticket, nq, ix->q_f32.as<float>(), k, id_base, ids_dev, sc_dev, st,
                                   nullptr, 0, packed_dev)));
    }
  }
  return B200_OK;
}

// move nrows rows of row_bytes each from src_row to dst_row (dst_row < src_row) inside one buffer
static int move_rows_down(uint8_t* base, size_t row_bytes, int64_t dst_row, int64_t src_row, int64_t nrows,
                          DevBuf& tmp, cudaStream_t st) {
  if (nrows <= 0 || dst_row == src_row) return B200_OK;
  const int64_t shift = src_row - dst_row;
  if (shift >= nrows) {
    B200_CUDA(cudaMemcpyAsync(base + dst_row * row_bytes, base + src_row * row_bytes, (size_t)nrows * row_bytes,
                              cudaMemcpyDeviceToDevice, st));
    return B200_OK;
  }
  const int64_t piece = std::max<int64_t>(shift, std::min<int64_t>(nrows, (int64_t)((64u << 20) / row_bytes)));
  if (piece > shift) B200_TRY(tmp.grow((size_t)piece * row_bytes, 0, st));
  for (int64_t o = 0; o < nrows; o += piece) {
    const int64_t m = std::min(piece, nrows - o);
    uint8_t* s = base + (src_row + o) * row_bytes;
    uint8_t* d = base + (dst_row + o) * row_bytes;
    if (m <= shift) {
      B200_CUDA(cudaMemcpyAsync(d, s, (size_t)m * row_bytes, cudaMemcpyDeviceToDevice, st));
    } else {
      B200_CUDA(cudaMemcpyAsync(tmp.p, s, (size_t)m * row_bytes, cudaMemcpyDeviceToDevice, st));
      B200_CUDA(cudaMemcpyAsync(d, tmp.p, (size_t)m * row_bytes, cudaMemcpyDeviceToDevice, st));
    }
  }
  return B200_OK;
}

extern "C" {

const char* b200_last_error(void) { return b200::get_error(); }
int b200_version(void) { return 100; }

int b200_set_option(const char* name, int value) {
  if (!name) B200_FAIL(B200_ERR_INVALID, "option name is NULL");
  opts_init();
  for (int i = 0; i < OPT_COUNT; ++i) {
    if (strcmp(name, kOptDefs[i].name) != 0) continue;
    if (value < kOptDefs[i].lo || value > kOptDefs[i].hi)
      B200_FAIL(B200_ERR_INVALID, "option '%s' takes %d..%d, got %d", name, kOptDefs[i].lo, kOptDefs[i].hi, value);
    g_opts[i].store(value, std::memory_order_relaxed);
    return B200_OK;
  }
  B200_FAIL(B200_ERR_INVALID, "unknown option '%s'", name);
}

int b200_get_option(const char* name, int* value) {
  if (!name || !value) B200_FAIL(B200_ERR_INVALID, "NULL argument");
  opts_init();
  for (int i = 0; i < OPT_COUNT; ++i) {
    if (strcmp(name, kOptDefs[i].name) != 0) continue;
    *value = g_opts[i].load(std::memory_order_relaxed);
    return B200_OK;
  }
  B200_FAIL(B200_ERR_INVALID, "unknown option '%s'", name);
}

int b200_profile_enable(int on) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_on = on != 0;
  return B200_OK;
}

int b200_profile_reset(void) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (auto& r : g_prof) {
    cudaEventDestroy(r.a);
    cudaEventDestroy(r.b);
  }
  g_prof.clear();
  return B200_OK;
}

int64_t b200_profile_read(char* buf, int64_t cap) {
  if (!buf || cap <= 0) B200_FAIL(B200_ERR_INVALID, "bad profile buffer");
  if (cudaDeviceSynchronize() != cudaSuccess) B200_FAIL(B200_ERR_CUDA, "cudaDeviceSynchronize failed in profile_read");
  std::lock_guard<std::mutex> lk(g_prof_mu);
  struct Agg { const char* name; long n; double ms, flops, bytes; };
  std::vector<Agg> agg;
  for (auto& r : g_prof) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.a, r.b) != cudaSuccess) { cudaGetLastError(); continue; }
    Agg* a = nullptr;
    for (auto& x : agg) if (strcmp(x.name, r.name) == 0) a = &x;
    if (!a) { agg.push_back({r.name, 0, 0, 0, 0}); a = &agg.back(); }
    a->n += 1; a->ms += ms; a->flops += r.flops; a->bytes += r.bytes;
  }
  int64_t off = 0;
  for (auto& a : agg) {
    int w = snprintf(buf + off, (size_t)(cap - off), "%s %ld %.6f %.6e %.6e\n", a.name, a.n, a.ms, a.flops, a.bytes);
    if (w < 0 || off + w >= cap) break;
    off += w;
  }
  buf[std::min(off, cap - 1)] = 0;
  return off;
}

int b200_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  int ok = 0;
  for (int d = 0; d < n; ++d) {
    int major = 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, d) == cudaSuccess && major == 10) ++ok;
  }
  return ok;
}

int b200_index_create(int dim, int64_t capacity, int device, int master_dtype, b200_index_t** out) {
  if (!out) B200_FAIL(B200_ERR_INVALID, "out is NULL");
  *out = nullptr;
  if (dim <= 0) B200_FAIL(B200_ERR_INVALID, "dim must be positive, got %d", dim);
  if (dim > 4096) B200_FAIL(B200_ERR_UNSUPPORTED, "dim %d > 4096 is not supported", dim);
  if (master_dtype != B200_INDEX_MASTER_F32 && master_dtype != B200_INDEX_MASTER_BF16)
    B200_FAIL(B200_ERR_INVALID, "unknown master_dtype %d", master_dtype);
  B200_TRY(check_device_is_sm100(device));
  DeviceGuard g(device);
  if (!g.ok) B200_FAIL(B200_ERR_CUDA, "cudaSetDevice(%d) failed", device);
  b200_index* ix = new b200_index();
  ix->dim = dim;
  ix->dpad = (dim + 63) / 64 * 64;
  ix->device = device;
  ix->master_dtype = master_dtype;
  cudaDeviceGetAttribute(&ix->num_sms, cudaDevAttrMultiProcessorCount, device);
  int rc = index_reserve_locked(ix, std::max<int64_t>(capacity, 1024), 0);
  if (rc == B200_OK) rc = ix->persist.alloc(64);
  if (rc == B200_OK && cudaMemset(ix->persist.p, 0, 64) != cudaSuccess) { set_error("cudaMemset failed"); rc = B200_ERR_CUDA; }
  if (rc == B200_OK) rc = ix->order.init();
  if (rc != B200_OK) {
    delete ix;
    return rc;
  }
  *out = ix;
  return B200_OK;
}

int b200_index_destroy(b200_index_t* ix) {
  if (!ix) return B200_OK;
  {
    DeviceGuard g(ix->device);
    cudaDeviceSynchronize();
    delete ix;
  }
  return B200_OK;
}

int b200_index_clear(b200_index_t* ix) {
  if (!ix) B200_FAIL(B200_ERR_INVALID, "index is NULL");
  std::lock_guard<std::mutex> lk(ix->mu);
  ix->count = 0;
  return B200_OK;
}

int64_t b200_index_count(const b200_index_t* ix) { return ix ? ix->count : 0; }
int b200_index_dim(const b200_index_t* ix) { return ix ? ix->dim : 0; }

int b200_index_reserve(b200_index_t* ix, int64_t capacity) {
  if (!ix) B200_FAIL(B200_ERR_INVALID, "index is NULL");
  std::lock_guard<std::mutex> lk(ix->mu);
  DeviceGuard g(ix->device);
  return index_reserve_locked(ix, capacity, 0);
}

int b200_index_add(b200_index_t* ix, const void* rows, int64_t n, int rows_on_device, void* stream) {
  if (!ix) B200_FAIL(B200_ERR_INVALID, "index is NULL");
  if (n < 0 || (n > 0 && !rows)) B200_FAIL(B200_ERR_INVALID, "bad rows pointer / count");
  if (n == 0) return B200_OK;
  std::lock_guard<std::mutex> lk(ix->mu);
  DeviceGuard g(ix->device);
  cudaStream_t st = (cudaStream_t)stream;
  B200_TRY(ix->order.enter(st));
  B200_TRY(index_reserve_locked(ix, ix->count + n, st));
  const size_t elt = ix->master_elt();
  const void* src = rows;
  if (!rows_on_device) {
    B200_TRY(ix->stage.grow((size_t)n * ix->dim * elt, 0, st));
    B200_CUDA(cudaMemcpyAsync(ix->stage.p, rows, (size_t)n * ix->dim * elt, cudaMemcpyHostToDevice, st));
    src = ix->stage.p;
  }
  const int wpb = 8;
  const unsigned grid = (unsigned)((n + wpb - 1) / wpb);
  if (ix->master_dtype == B200_INDEX_MASTER_F32) {
    if (ix->dim == ix->dpad && (ix->dim % 8) == 0 && ((uintptr_t)src % 16) == 0)
      ingest_rows_f32_vec_kernel<<<grid, wpb * 32, 0, st>>>((const float*)src, n, ix->dim, ix->search.as<__half>(),
                                                           ix->count, ix->master.as<float>(), ix->rvmax());
    else
      ingest_rows_kernel<float><<<grid, wpb * 32, 0, st>>>((const float*)src, n, ix->dim, ix->dpad, ix->search.as<__half>(),
                                                           ix->count, ix->master.as<float>(), ix->rvmax());
  } else {
    ingest_rows_kernel<__nv_bfloat16><<<grid, wpb * 32, 0, st>>>((const __nv_bfloat16*)src, n, ix->dim, ix->dpad,
                                                                  ix->search.as<__half>(), ix->count,
                                                                  ix->master.as<__nv_bfloat16>(), ix->rvmax());
  }
  B200_CUDA(cudaGetLastError());
  B200_TRY(ix->order.leave(st));
  if (!rows_on_device) B200_CUDA(cudaStreamSynchronize(st));  // the staging buffer and the caller's rows are reusable
  ix->count += n;
  return B200_OK;
}

int b200_index_remove_rows(b200_index_t* ix, const int64_t* rows, int64_t n, void* stream) {
  if (!ix) B200_FAIL(B200_ERR_INVALID, "index is NULL");
  if (n < 0 || (n > 0 && !rows)) B200_FAIL(B200_ERR_INVALID, "bad rows pointer / count");
  if (n == 0) return B200_OK;
  std::lock_guard<std::mutex> lk(ix->mu);
  for (int64_t i = 0; i < n; ++i) {
    if (rows[i] < 0 || rows[i] >= ix->count) B200_FAIL(B200_ERR_NOT_FOUND, "row %lld not found (count %lld)", (long long)rows[i], (long long)ix->count);
    if (i > 0 && rows[i] <= rows[i - 1]) B200_FAIL(B200_ERR_INVALID, "rows to remove must be ascending and unique");
  }
  DeviceGuard g(ix->device);
  cudaStream_t st = (cudaStream_t)stream;
  B200_TRY(ix->order.enter(st));
  const size_t sb = (size_t)ix->dpad * 2, mb = (size_t)ix->dim * ix->master_elt();
  int64_t dst = rows[0];
  for (int64_t i = 0; i < n; ++i) {
    const int64_t seg0 = rows[i] + 1;
    const int64_t seg1 = (i + 1 < n) ? rows[i + 1] : ix->count;
    B200_TRY(move_rows_down(ix->search.as<uint8_t>(), sb, dst, seg0, seg1 - seg0, ix->tmp, st));
    B200_TRY(move_rows_down(ix->master.as<uint8_t>(), mb, dst, seg0, seg1 - seg0, ix->tmp, st));
    dst += seg1 - seg0;
  }
  B200_TRY(ix->order.leave(st));
  ix->count -= n;
  return B200_OK;
}

int b200_index_get_rows(b200_index_t* ix, const int64_t* rows, int64_t n, float* out, int out_on_device, void* stream) {
  if (!ix) B200_FAIL(B200_ERR_INVALID, "index is NULL");
  if (n < 0 || (n > 0 && (!rows || !out))) B200_FAIL(B200_ERR_INVALID, "bad pointer / count");
  if (n == 0) return B200_OK;
  std::lock_guard<std::mutex> lk(ix->mu);
  for (int64_t i = 0; i < n; ++i)
    if (rows[i] < 0 || rows[i] >= ix->count) B200_FAIL(B200_ERR_NOT_FOUND, "row %lld not found (count %lld)", (long long)rows[i], (long long)ix->count);
  DeviceGuard g(ix->device);
  cudaStream_t st = (cudaStream_t)stream;
  B200_TRY(ix->order.enter(st));
  const size_t idx_bytes = ((size_t)n * 8 + 255) / 256 * 256;
  const size_t out_bytes = (size_t)n * ix->dim * 4;
  B200_TRY(ix->stage.grow(idx_bytes + (out_on_device ? 0 : out_bytes), 0, st));
  int64_t* d_rows = ix->stage.as<int64_t>();
  float* d_out = out_on_device ? out : reinterpret_cast<float*>(ix->stage.as<uint8_t>() + idx_bytes);
  B200_CUDA(cudaMemcpyAsync(d_rows, rows, (size_t)n * 8, cudaMemcpyHostToDevice, st));
  if (ix->master_dtype == B200_INDEX_MASTER_F32)
    gather_rows_kernel<float, float><<<(unsigned)n, 128, 0, st>>>(ix->master.as<float>(), d_rows, n, ix->dim, d_out);
  else
    gather_rows_kernel<__nv_bfloat16, float><<<(unsigned)n, 128, 0, st>>>(ix->master.as<__nv_bfloat16>(), d_rows, n, ix->dim, d_out);
  B200_CUDA(cudaGetLastError());
  if (!out_on_device) B200_CUDA(cudaMemcpyAsync(out, d_out, out_bytes, cudaMemcpyDeviceToHost, st));
  B200_TRY(ix->order.leave(st));
  B200_CUDA(cudaStreamSynchronize(st));
  return B200_OK;
}

}  // extern "C"

// top_k in (48, EXH_KEEP]: exhaustive exact scoring, one query at a time (see search_kernels.cuh).
static int search_exhaustive(b200_index* ix, const uint8_t* q_dev, int q_dtype, int64_t nq, int k, int64_t id_base,
                             int64_t* d_ids, float* d_sc, HitSink d_packed, cudaStream_t st) {
  const uint32_t n_rows = (uint32_t)ix->count;
  const uint64_t chunks0 = ((uint64_t)n_rows + EXH_CHUNK - 1) / EXH_CHUNK;
  // scratch: fp32 copy of one query, keys of every row, two ping-pong survivor buffers
  const size_t q_bytes = ((size_t)ix->dim * 4 + 255) / 256 * 256;
  const size_t keys_bytes = (size_t)n_rows * 8;
  const size_t surv_bytes = (size_t)chunks0 * EXH_KEEP * 8;
  B200_TRY(ix->tmp.grow(q_bytes + keys_bytes + 2 * surv_bytes + 512, 0, st));
  float* q_f32 = ix->tmp.as<float>();
  uint64_t* keys = reinterpret_cast<uint64_t*>(ix->tmp.as<uint8_t>() + q_bytes);
  uint64_t* surv[2] = {keys + n_rows, keys + n_rows + chunks0 * EXH_KEEP};
  B200_CUDA(cudaFuncSetAttribute(exh_reduce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, EXH_CHUNK * 8));
  const int score_grid = (int)std::min<uint64_t>((n_rows + 7) / 8, (uint64_t)ix->num_sms * 8);
  for (int64_t q = 0; q < nq; ++q) {
    if (q_dtype == 0) {
      B200_CUDA(cudaMemcpyAsync(q_f32, q_dev + (size_t)q * ix->dim * 4, (size_t)ix->dim * 4, cudaMemcpyDeviceToDevice, st));
    } else {
      exh_query_f32_kernel<<<1, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(q_dev) + (size_t)q * ix->dim, ix->dim, q_f32);
    }
    if (ix->master_dtype == B200_INDEX_MASTER_F32)
      exh_score_kernel<float><<<score_grid, 256, (size_t)ix->dim * 4, st>>>(q_f32, ix->master.as<float>(), n_rows, ix->dim, keys);
    else
      exh_score_kernel<__nv_bfloat16><<<score_grid, 256, (size_t)ix->dim * 4, st>>>(q_f32, ix->master.as<__nv_bfloat16>(), n_rows,
                                                                                     ix->dim, keys);
    const uint64_t* in = keys;
    uint64_t n_in = n_rows;
    int pp = 0;
    for (;;) {
      const uint64_t chunks = (n_in + EXH_CHUNK - 1) / EXH_CHUNK;
      exh_reduce_kernel<<<(unsigned)chunks, 512, EXH_CHUNK * 8, st>>>(in, n_in, surv[pp]);
      in = surv[pp];
      n_in = chunks * EXH_KEEP;
      pp ^= 1;
      if (chunks == 1) break;
    }
    exh_write_kernel<<<1, 256, 0, st>>>(in, EXH_KEEP, k, id_base, sink_on(d_packed) ? nullptr : d_ids + q * k,
                                        sink_on(d_packed) ? nullptr : d_sc + q * k, d_packed, sink_on(d_packed) ? (size_t)(q * k) : 0);
    B200_CUDA(cudaGetLastError());
  }
  return B200_OK;
}

// b200_index_search and b200_index_search_packed: `packed` (device, [nq][k] 16-byte records) replaces ids / scores
static int search_impl(b200_index* ix, const void* queries, int q_dtype, int queries_on_device, int64_t nq, int k,
                       int64_t id_base, int64_t* ids, float* scores, HitSink packed, int out_on_device, void* stream) {
  const bool is_packed = sink_on(packed);
  if (!ix) B200_FAIL(B200_ERR_INVALID, "index is NULL");
  if (nq < 0 || k < 0) B200_FAIL(B200_ERR_INVALID, "negative nq / k");
  if (nq == 0 || k == 0) return B200_OK;
  if (!queries || (!is_packed && (!ids || !scores))) B200_FAIL(B200_ERR_INVALID, "NULL buffer");
  if (q_dtype != 0 && q_dtype != 1) B200_FAIL(B200_ERR_INVALID, "q_dtype must be 0 (fp32) or 1 (bf16)");
  if (k > EXH_KEEP) B200_FAIL(B200_ERR_UNSUPPORTED, "k=%d: top_k above %d is not supported", k, EXH_KEEP);
  const int kp = k <= 10 ? kKpSmall : (k <= 24 ? 32 : 64);
  std::lock_guard<std::mutex> lk(ix->mu);
  DeviceGuard g(ix->device);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t qelt = q_dtype == 0 ? 4 : 2;
  B200_TRY(ix->order.enter(st));

  int64_t* d_ids = ids;
  float* d_sc = scores;
  if (!is_packed && !out_on_device) {
    B200_TRY(ix->out_ids.grow((size_t)nq * k * 8, 0, st));
    B200_TRY(ix->out_scores.grow((size_t)nq * k * 4, 0, st));
    d_ids = ix->out_ids.as<int64_t>();
    d_sc = ix->out_scores.as<float>();
  }
  if (ix->count == 0) {  // inmemory.py:93-94: empty store -> no results
    const size_t n = (size_t)nq * k;
    fill_no_hits_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 1024), 256, 0, st>>>(d_ids, d_sc, packed, n);
    B200_CUDA(cudaGetLastError());
  } else {
    const uint8_t* q_dev = (const uint8_t*)queries;
    if (!queries_on_device) {
      B200_TRY(ix->stage.grow((size_t)nq * ix->dim * qelt, 0, st));
      B200_CUDA(cudaMemcpyAsync(ix->stage.p, queries, (size_t)nq * ix->dim * qelt, cudaMemcpyHostToDevice, st));
      q_dev = ix->stage.as<uint8_t>();
    }
    if (k > 48) {
      B200_TRY(search_exhaustive(ix, q_dev, q_dtype, nq, k, id_base, d_ids, d_sc, packed, st));
    } else {
      for (int64_t q0 = 0; q0 < nq; q0 += kMaxQueryChunk) {
        const int n = (int)std::min<int64_t>(kMaxQueryChunk, nq - q0);
        B200_TRY(search_chunk(ix, q_dev + (size_t)q0 * ix->dim * qelt, q_dtype, n, k, kp, id_base,
                              is_packed ? nullptr : d_ids + q0 * k, is_packed ? nullptr : d_sc + q0 * k,
                              is_packed ? sink_at(packed, (size_t)(q0 * k)) : HitSink{}, st));
      }
    }
  }
  if (!is_packed && !out_on_device) {
    B200_CUDA(cudaMemcpyAsync(ids, d_ids, (size_t)nq * k * 8, cudaMemcpyDeviceToHost, st));
    B200_CUDA(cudaMemcpyAsync(scores, d_sc, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, st));
  }
  B200_TRY(ix->order.leave(st));
  if (!is_packed && !out_on_device) B200_CUDA(cudaStreamSynchronize(st));
  return B200_OK;
}

extern "C" {

int b200_index_search_exchange(b200_index_t* ix, const void* queries_dev, int q_dtype, int64_t nq, int k, int64_t id_base,
                               void* const* peer_hits, void* const* peer_flags, int world, int rank, uint32_t epoch,
                               void* stream) {
  if (!ix) B200_FAIL(B200_ERR_INVALID, "index is NULL");
  if (!peer_hits || !peer_flags) B200_FAIL(B200_ERR_INVALID, "NULL buffer");
  if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world)
    B200_FAIL(B200_ERR_UNSUPPORTED, "peer exchange covers 1..%d ranks (got world %d, rank %d)", kMaxPeers, world, rank);
  HitSink sink{}, flags{};
  sink.n_peers = flags.n_peers = world;
  for (int i = 0; i < world; ++i) {
    if (!peer_hits[i] || !peer_flags[i]) B200_FAIL(B200_ERR_INVALID, "NULL peer buffer %d", i);
    sink.peer[i] = static_cast<ulonglong2*>(peer_hits[i]);
    flags.peer[i] = static_cast<ulonglong2*>(peer_flags[i]);
  }
  if (nq > 0 && k > 0) B200_TRY(search_impl(ix, queries_dev, q_dtype, 1, nq, k, id_base, nullptr, nullptr, sink, 1, stream));
  DeviceGuard g(ix->device);
  signal_peers_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(flags, rank, epoch);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

int b200_index_search(b200_index_t* ix, const void* queries, int q_dtype, int queries_on_device, int64_t nq, int k,
                      int64_t id_base, int64_t* ids, float* scores, int out_on_device, void* stream) {
  return search_impl(ix, queries, q_dtype, queries_on_device, nq, k, id_base, ids, scores, HitSink{}, out_on_device, stream);
}

int b200_index_search_packed(b200_index_t* ix, const void* queries_dev, int q_dtype, int64_t nq, int k, int64_t id_base,
                             void* hits_dev, void* stream) {
  if (!hits_dev) B200_FAIL(B200_ERR_INVALID, "NULL buffer");
  HitSink sink{};
  sink.one = static_cast<ulonglong2*>(hits_dev);
  return search_impl(ix, queries_dev, q_dtype, 1, nq, k, id_base, nullptr, nullptr, sink, 1, stream);
}

static int merge_hits_impl(const void* hits_in, int nlists, int64_t nq_per_list, int k_in, int64_t q_first, int64_t nq, int k_out,
                           int64_t* ids_out, float* scores_out, const uint32_t* wait_flags, uint32_t epoch, int device,
                           void* stream) {
  if (nq < 0 || nlists <= 0 || k_in <= 0 || k_out <= 0 || q_first < 0 || q_first + nq > nq_per_list)
    B200_FAIL(B200_ERR_INVALID, "bad merge shape");
  if (nq == 0) return B200_OK;
  if (!hits_in || !ids_out || !scores_out) B200_FAIL(B200_ERR_INVALID, "NULL buffer");
  if ((int64_t)nlists * k_in > 4096) B200_FAIL(B200_ERR_UNSUPPORTED, "nlists*k_in = %lld > 4096", (long long)nlists * k_in);
  DeviceGuard g(device);
  if (!g.ok) B200_FAIL(B200_ERR_CUDA, "cudaSetDevice(%d) failed", device);
  const int wpb = 4;
  {
    B200_PROF("merge_hits", 0.0, (double)nq * nlists * k_in * 16.0, (cudaStream_t)stream);
    merge_packed_kernel<<<(unsigned)((nq + wpb - 1) / wpb), wpb * 32, 0, (cudaStream_t)stream>>>(
        static_cast<const ulonglong2*>(hits_in), nlists, (int)nq_per_list, k_in, (int)q_first, (int)nq, k_out, ids_out, scores_out,
        wait_flags, epoch);
  }
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

int b200_merge_hits(const void* hits_in, int nlists, int64_t nq_per_list, int k_in, int64_t q_first, int64_t nq, int k_out,
                    int64_t* ids_out, float* scores_out, int device, void* stream) {
  return merge_hits_impl(hits_in, nlists, nq_per_list, k_in, q_first, nq, k_out, ids_out, scores_out, nullptr, 0, device, stream);
}

int b200_merge_hits_wait(const void* hits_local, const void* flags_local, uint32_t epoch, int nlists, int64_t nq_per_list, int k_in,
                         int64_t q_first, int64_t nq, int k_out, int64_t* ids_out, float* scores_out, int device, void* stream) {
  if (!flags_local) B200_FAIL(B200_ERR_INVALID, "NULL flags");
  if (nlists > 32) B200_FAIL(B200_ERR_UNSUPPORTED, "at most 32 lists can be waited for");
  return merge_hits_impl(hits_local, nlists, nq_per_list, k_in, q_first, nq, k_out, ids_out, scores_out,
                         static_cast<const uint32_t*>(flags_local), epoch, device, stream);
}

int b200_index_certificate_stats(b200_index_t* ix, int64_t* checked, int64_t* recomputed) {
  if (!ix || !checked || !recomputed) B200_FAIL(B200_ERR_INVALID, "NULL argument");
  std::lock_guard<std::mutex> lk(ix->mu);
  DeviceGuard g(ix->device);
  unsigned long long h[2] = {0, 0};
  B200_CUDA(cudaDeviceSynchronize());
  B200_CUDA(cudaMemcpy(h, ix->stats(), sizeof(h), cudaMemcpyDeviceToHost));
  *checked = (int64_t)h[0];
  *recomputed = (int64_t)h[1];
  return B200_OK;
}

int b200_index_search_l2(b200_index_t* ix, const void* queries, int q_dtype, int queries_on_device, int64_t nq, int k,
                         int64_t id_base, int64_t* ids, float* dist, int out_on_device, void* stream) {
  if (!ix) B200_FAIL(B200_ERR_INVALID, "index is NULL");
  if (nq < 0 || k < 0) B200_FAIL(B200_ERR_INVALID, "negative nq / k");
  if (nq == 0 || k == 0) return B200_OK;
  if (!queries || !ids || !dist) B200_FAIL(B200_ERR_INVALID, "NULL buffer");
  if (q_dtype != 0) B200_FAIL(B200_ERR_INVALID, "L2 search takes fp32 queries (q_dtype 0)");
  if (k > EXACT_KEEP) B200_FAIL(B200_ERR_UNSUPPORTED, "k=%d: L2 search returns at most %d neighbours per query", k, EXACT_KEEP);
  std::lock_guard<std::mutex> lk(ix->mu);
  DeviceGuard g(ix->device);
  cudaStream_t st = (cudaStream_t)stream;
  B200_TRY(ix->order.enter(st));
  int64_t* d_ids = ids;
  float* d_sc = dist;
  if (!out_on_device) {
    B200_TRY(ix->out_ids.grow((size_t)nq * k * 8, 0, st));
    B200_TRY(ix->out_scores.grow((size_t)nq * k * 4, 0, st));
    d_ids = ix->out_ids.as<int64_t>();
    d_sc = ix->out_scores.as<float>();
  }
  if (ix->count == 0) {
    B200_CUDA(cudaMemsetAsync(d_ids, 0xFF, (size_t)nq * k * 8, st));
    B200_CUDA(cudaMemsetAsync(d_sc, 0, (size_t)nq * k * 4, st));
  } else {
    const float* q_dev = (const float*)queries;
    if (!queries_on_device) {
      B200_TRY(ix->stage.grow((size_t)nq * ix->dim * 4, 0, st));
      B200_CUDA(cudaMemcpyAsync(ix->stage.p, queries, (size_t)nq * ix->dim * 4, cudaMemcpyHostToDevice, st));
      q_dev = ix->stage.as<float>();
    }
    for (int64_t q0 = 0; q0 < nq; q0 += kMaxQueryChunk) {
      const int n = (int)std::min<int64_t>(kMaxQueryChunk, nq - q0);
      B200_TRY(ix->cert.grow(64 + (size_t)n * 8, 0, st));
      B200_CUDA(cudaMemsetAsync(ix->cert.p, 0, 64 + (size_t)n * 4, st));
      B200_TRY((launch_exact<1, 0>(ix, nullptr, nullptr, ix->cert.as<uint32_t>() + 16, n, q_dev + (size_t)q0 * ix->dim, k, id_base,
                               d_ids + q0 * k, d_sc + q0 * k, st)));
    }
  }
  if (!out_on_device) {
    B200_CUDA(cudaMemcpyAsync(ids, d_ids, (size_t)nq * k * 8, cudaMemcpyDeviceToHost, st));
    B200_CUDA(cudaMemcpyAsync(dist, d_sc, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, st));
  }
  B200_TRY(ix->order.leave(st));
  if (!out_on_device) B200_CUDA(cudaStreamSynchronize(st));
  return B200_OK;
}

int b200_merge_topk(const int64_t* ids_in, const float* scores_in, int64_t nq, int nlists, int k_in, int k_out,
                    int64_t* ids_out, float* scores_out, int device, void* stream) {
  if (nq < 0 || nlists <= 0 || k_in <= 0 || k_out <= 0) B200_FAIL(B200_ERR_INVALID, "bad merge shape");
  if (nq == 0) return B200_OK;
  if (!ids_in || !scores_in || !ids_out || !scores_out) B200_FAIL(B200_ERR_INVALID, "NULL buffer");
  if ((int64_t)nlists * k_in > 4096) B200_FAIL(B200_ERR_UNSUPPORTED, "nlists*k_in = %lld > 4096", (long long)nlists * k_in);
  DeviceGuard g(device);
  if (!g.ok) B200_FAIL(B200_ERR_CUDA, "cudaSetDevice(%d) failed", device);
  const int wpb = 4;
  merge_lists_kernel<<<(unsigned)((nq + wpb - 1) / wpb), wpb * 32, 0, (cudaStream_t)stream>>>(
      ids_in, scores_in, (int)nq, nlists, k_in, k_out, ids_out, scores_out);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

}  // extern "C"

// ================================================================================================
// Search that never allocates.
//
// Some programs that embed this library cannot let it call cudaMalloc on the query path: a serving
// process that handed every free byte of HBM to its own memory pool at start-up (a later cudaMalloc
// from a third-party library simply fails there), and latency-critical query threads for which the
// device-wide synchronisation hidden inside cudaMalloc / cudaFree would stall every other stream.
// b200_index_search() grows its scratch buffers on demand, so it is off limits for them.
//
// b200_index_search_static() does the same search with all of its scratch -- the padded bf16 query
// tile, the fp32 copy of the queries, and the per-split candidate lists -- inside the fixed array
// below, which is part of the library image: every GPU that loads the library holds these 4.7 MB
// until the process exits, whether or not the entry point is ever used.
// ================================================================================================
namespace {

constexpr int kStaticMaxQueries = SIM_BLOCK_M;      // one query tile
constexpr int kStaticMaxDpad = 384;
constexpr int kStaticMaxCand = 4096;                // splits x kept candidates, per query
constexpr size_t kStaticOffQBf16 = 0;
constexpr size_t kStaticOffQF32 = kStaticOffQBf16 + (size_t)kStaticMaxQueries * kStaticMaxDpad * 2;
constexpr size_t kStaticOffCand = kStaticOffQF32 + (size_t)kStaticMaxQueries * kStaticMaxDpad * 4;
constexpr size_t kStaticOffGthr = kStaticOffCand + (size_t)kStaticMaxQueries * kStaticMaxCand * 8;
constexpr size_t kStaticBktBytes = (size_t)kStaticMaxQueries * (64 + 1) * 4;   // [128][up to 64 buckets] + [128] bounds
// the exactness certificate's share: per-query rounding data, the fail list, and the exhaustive kernel's lists
constexpr int kStaticExactGrid = 160;               // CTAs of the exhaustive kernel (at most one per SM)
constexpr size_t kStaticOffQMeta = kStaticOffGthr + (kStaticBktBytes + 255) / 256 * 256;
constexpr size_t kStaticOffCert = kStaticOffQMeta + (size_t)kStaticMaxQueries * 8;
constexpr size_t kStaticCertBytes = 64 + (size_t)kStaticMaxQueries * 8;         // count, tickets, fail list
constexpr size_t kStaticOffPart = kStaticOffCert + (kStaticCertBytes + 255) / 256 * 256;
constexpr size_t kStaticBytes = kStaticOffPart + (size_t)2 * kStaticExactGrid * EXACT_KEEP * 8;
static_assert(kStaticOffQF32 % 256 == 0 && kStaticOffCand % 256 == 0 && kStaticOffGthr % 256 == 0 &&
              kStaticOffQMeta % 256 == 0 && kStaticOffCert % 256 == 0 && kStaticOffPart % 256 == 0,
              "scratch sub-buffers must stay 256-byte aligned");

}  // namespace

static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

namespace {

// In host code `g_static_scratch` is only a handle for the symbol, not a device address: the real
// address differs per device and has to be asked for with that device current.
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
unsigned char* g_static_ptr[kMaxDevices] = {};

int static_scratch_for_current_device(int device, unsigned char** out) {
  if (device < 0 || device >= kMaxDevices) B200_FAIL(B200_ERR_INVALID, "device %d out of range", device);
  if (!g_static_ptr[device]) {
    void* p = nullptr;
    B200_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
    // With lazy module loading (the CUDA 12 default) a kernel's code is loaded into device memory the first time
    // it is used.  Touch every kernel this entry point can launch now, so that a later call in another k band or
    // query dtype loads nothing.
    cudaFuncAttributes fa;
    B200_CUDA(cudaFuncGetAttributes(&fa, sim_topk_kernel<256, kKpSmall, 1>));
    B200_CUDA(cudaFuncGetAttributes(&fa, sim_topk_kernel<128, 32, 1>));
    B200_CUDA(cudaFuncGetAttributes(&fa, sim_topk_kernel<128, 64, 1>));
    B200_CUDA(cudaFuncGetAttributes(&fa, prep_queries_kernel<float>));
    B200_CUDA(cudaFuncGetAttributes(&fa, prep_queries_kernel<__nv_bfloat16>));
    B200_CUDA(cudaFuncGetAttributes(&fa, merge_rescore_kernel));
    B200_CUDA(cudaFuncGetAttributes(&fa, exact_topk_kernel<0, 0>));
    g_static_ptr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_static_ptr[device];
  return B200_OK;
}

template <int BLOCK_N, int KP>
int launch_sim_static(b200_index* ix, int nq, const CUtensorMap& map_q, uint64_t* cand, uint32_t* gthr,
                      cudaStream_t st, SimParams* used) {
  const int kc = ix->dpad / SIM_BLOCK_K;
  const size_t a_bytes = (size_t)kc * SIM_A_CHUNK_BYTES;
  const size_t list_bytes = (size_t)SIM_BLOCK_M * (KP + 1) * 8;
  const size_t stage_bytes = (size_t)BLOCK_N * SIM_BLOCK_K * 2;
  const size_t fixed = 1024 + a_bytes + list_bytes + 256;
  const size_t max_smem = 232448;
  const int nstage = (int)std::min<size_t>(8, (max_smem - fixed) / stage_bytes);
  const size_t smem = fixed + (size_t)nstage * stage_bytes;
  const uint32_t n_rows = (uint32_t)ix->count;
  const uint32_t total_tiles = (n_rows + BLOCK_N - 1) / BLOCK_N;
  // one wave of CTAs, bounded by what the fixed candidate area can take
  int nsplit = std::max(1, std::min<int>({(int)std::min<uint32_t>(total_tiles, 65535u), kStaticMaxCand / KP, ix->num_sms}));
  const uint32_t tps = (total_tiles + nsplit - 1) / nsplit;
  nsplit = (int)((total_tiles + tps - 1) / tps);
  CUtensorMap map_c;
  B200_TRY(make_tma_2d_bf16(&map_c, ix->search.p, (uint64_t)ix->count, (uint64_t)ix->dpad, BLOCK_N, SIM_BLOCK_K));
  SimParams p{};
  p.nq = nq; p.kc = kc; p.nstage = nstage; p.n_rows = n_rows; p.tiles_per_split = tps; p.nsplit = nsplit;
  p.bkt_per_cta = nsplit >= KP ? 1 : (KP + nsplit - 1) / nsplit;
  p.cand = cand; p.gbkt = gthr; p.use_bound = (int)tps >= p.bkt_per_cta;
  p.gthr = gthr + (size_t)kStaticMaxQueries * KP;
  p.refresh_mask = 0; p.use_gthr = 0;
  auto kern = sim_topk_kernel<BLOCK_N, KP, 1>;
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<dim3(1, nsplit), SIM_THREADS, smem, st>>>(map_q, map_c, p);
  B200_CUDA(cudaGetLastError());
  *used = p;
  return B200_OK;
}

}  // namespace

extern "C" int b200_index_search_static(b200_index_t* ix, const void* queries_dev, int q_dtype, int64_t nq, int k,
                                        int64_t id_base, int64_t* ids_dev, float* scores_dev, void* stream) {
  if (!ix) B200_FAIL(B200_ERR_INVALID, "index is NULL");
  if (nq < 0 || k < 0) B200_FAIL(B200_ERR_INVALID, "negative nq / k");
  if (nq == 0 || k == 0) return B200_OK;
  if (!queries_dev || !ids_dev || !scores_dev) B200_FAIL(B200_ERR_INVALID, "NULL buffer");
  if (q_dtype != 0 && q_dtype != 1) B200_FAIL(B200_ERR_INVALID, "q_dtype must be 0 (fp32) or 1 (bf16)");
  if (nq > kStaticMaxQueries)
    B200_FAIL(B200_ERR_UNSUPPORTED, "b200_index_search_static takes at most %d queries per call (got %lld): its scratch "
              "area is fixed; split the batch or use b200_index_search", kStaticMaxQueries, (long long)nq);
  if (ix->dpad > kStaticMaxDpad)
    B200_FAIL(B200_ERR_UNSUPPORTED, "b200_index_search_static covers dimensions up to %d (index has %d)", kStaticMaxDpad, ix->dim);
  if (k > 48) B200_FAIL(B200_ERR_UNSUPPORTED, "k=%d: the fused select keeps at most 64 candidates per query (k <= 48)", k);
  const int kp = k <= 10 ? kKpSmall : (k <= 24 ? 32 : 64);
  std::lock_guard<std::mutex> lk(ix->mu);
  std::lock_guard<std::mutex> lk2(g_static_mu);
  DeviceGuard g(ix->device);
  if (!g.ok) B200_FAIL(B200_ERR_CUDA, "cudaSetDevice(%d) failed", ix->device);
  cudaStream_t st = (cudaStream_t)stream;
  unsigned char* base = nullptr;
  B200_TRY(static_scratch_for_current_device(ix->device, &base));   // first call per device: resolves the symbol
  if (ix->count == 0) {  // inmemory.py:93-94: empty store -> no results
    B200_CUDA(cudaMemsetAsync(ids_dev, 0xFF, (size_t)nq * k * 8, st));
    B200_CUDA(cudaMemsetAsync(scores_dev, 0, (size_t)nq * k * 4, st));
    return B200_OK;
  }
  __half* q16 = reinterpret_cast<__half*>(base + kStaticOffQBf16);
  float* q_f32 = reinterpret_cast<float*>(base + kStaticOffQF32);
  uint64_t* cand = reinterpret_cast<uint64_t*>(base + kStaticOffCand);
  uint32_t* gthr = reinterpret_cast<uint32_t*>(base + kStaticOffGthr);
  float2* qmeta = reinterpret_cast<float2*>(base + kStaticOffQMeta);
  uint32_t* cert = reinterpret_cast<uint32_t*>(base + kStaticOffCert);
  uint64_t* part = reinterpret_cast<uint64_t*>(base + kStaticOffPart);
  // The array keeps whatever the previous call left.  What must start at zero is the table of score
  // bounds the CTAs share (33 KB) and the certificate's fail count and tickets (0.6 KB): they are cleared
  // here, on the caller's stream.  Everything else is rewritten in full before it is read:
  // prep_queries_kernel writes the whole padded query tile (zero rows included) and the rounding data of
  // every query, every (query, split) candidate set is stored by its CTA before the merge kernel looks at
  // it, and the exhaustive kernel writes its per-CTA lists before the CTA that folds them reads them.
  B200_CUDA(cudaMemsetAsync(gthr, 0, kStaticBktBytes, st));
  B200_CUDA(cudaMemsetAsync(cert, 0, 64 + (size_t)kStaticMaxQueries * 4, st));
  const int n = (int)nq, nq_pad = SIM_BLOCK_M, wpb = 8;
  if (q_dtype == 0)
    prep_queries_kernel<float><<<(nq_pad + wpb - 1) / wpb, wpb * 32, 0, st>>>((const float*)queries_dev, n, nq_pad, ix->dim,
                                                                           ix->dpad, q16, q_f32, qmeta);
  else
    prep_queries_kernel<__nv_bfloat16><<<(nq_pad + wpb - 1) / wpb, wpb * 32, 0, st>>>(
        (const __nv_bfloat16*)queries_dev, n, nq_pad, ix->dim, ix->dpad, q16, q_f32, qmeta);
  B200_CUDA(cudaGetLastError());
  CUtensorMap map_q;
  B200_TRY(make_tma_2d_bf16(&map_q, q16, (uint64_t)nq_pad, (uint64_t)ix->dpad, SIM_BLOCK_M, SIM_BLOCK_K));
  SimParams sp{};
  if (kp == kKpSmall) B200_TRY((launch_sim_static<256, kKpSmall>(ix, n, map_q, cand, gthr, st, &sp)));
  else if (kp == 32) B200_TRY((launch_sim_static<128, 32>(ix, n, map_q, cand, gthr, st, &sp)));
  else B200_TRY((launch_sim_static<128, 64>(ix, n, map_q, cand, gthr, st, &sp)));
  const int nsplit = sp.nsplit;
  MergeParams m;
  m.gbkt = sp.gbkt; m.gthr = sp.gthr; m.use_bound = sp.use_bound; m.use_gthr = sp.use_gthr;
  m.nq = n; m.nsplit = nsplit; m.kp = kp; m.k = k; m.d = ix->dim;
  int pow2 = 64;
  while (pow2 < nsplit * kp) pow2 <<= 1;
  m.ncand_pow2 = pow2;
  m.n_rows = (uint32_t)ix->count; m.id_base = id_base;
  m.cand = cand; m.q_f32 = q_f32; m.master = ix->master.p;
  m.master_is_bf16 = ix->master_dtype == B200_INDEX_MASTER_BF16;
  m.out_ids = ids_dev; m.out_scores = scores_dev; m.out_packed = HitSink{};
  // same exactness certificate as b200_index_search, its bookkeeping inside the static array
  uint32_t* ticket = cert + 16;
  int32_t* fail_idx = reinterpret_cast<int32_t*>(ticket + kStaticMaxQueries);
  m.qmeta = qmeta; m.rvmax = ix->rvmax(); m.slop = certificate_slop(ix->dpad);
  m.fail_count = cert; m.fail_idx = fail_idx; m.stats = ix->stats();
  const size_t msmem = (size_t)pow2 * 8 + (size_t)((ix->dim + 1) & ~1) * 4 + 64 * 8;
  B200_CUDA(cudaFuncSetAttribute(merge_rescore_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)msmem));
  merge_rescore_kernel<<<n, 256, msmem, st>>>(m);
  B200_CUDA(cudaGetLastError());
  B200_TRY(launch_exact<0>(ix, cert, fail_idx, ticket, n, q_f32, k, id_base, ids_dev, scores_dev, st, part,
                           std::min(ix->num_sms, kStaticExactGrid)));
  return B200_OK;
}

#ifdef B200_K8_TIMING
extern "C" int b200_debug_k8_timing(unsigned long long* out) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, b200::g_k8_timing, 128 * sizeof(unsigned long long));
  return 0;
}
#endif
