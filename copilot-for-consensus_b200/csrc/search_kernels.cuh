// search_kernels.cuh -- device code for the cosine top-k path.
//
//   ingest_rows_kernel   K11  raw rows -> unit-normalised 16-bit search copy (+ master copy)
//   prep_queries_kernel        raw queries -> zero-padded 16-bit tile operand + fp32 copy + rounding-error norm
//   sim_topk_kernel      K8   Q x C^T on tcgen05 (TMA-fed, TMEM accumulators) with the top-KP
//                             select fused into the TMEM drain: scores never reach HBM
//   merge_rescore_kernel K9   per query: merge the per-split candidate lists, rescore the KP
//                             survivors with the reference's fp32 arithmetic in the canonical
//                             summation order (oracle/search_oracle.c), emit top-k; certify that no
//                             row outside the survivors can belong to the exact top-k
//   exact_topk_kernel          queries that fail the certificate: exact score of EVERY row
//
// The search copy is IEEE half (E5M10), not bf16: its elements are components of unit vectors, which
// need mantissa, not range.  Same 2 bytes per element and the same tcgen05 kind::f16 rate, but an 8x
// smaller rounding error -- which is what makes the rigorous candidate-sufficiency bound of K9 tight
// enough to pass on all but a few queries in 10^5.  The caller's rows (bf16 or fp32) are kept untouched
// in the master copy.
//
// Reference semantics being reproduced: InMemoryVectorStore.query,
// adapters/copilot_vectorstore/copilot_vectorstore/inmemory.py:96-119.
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"

namespace b200 {

constexpr int SIM_BLOCK_M = 128;   // queries per CTA = TMEM lanes
constexpr int SIM_BLOCK_K = 64;    // 16-bit elements per 128-byte swizzle row
constexpr int SIM_THREADS = 256;   // warp 0 TMA, 1 MMA, 2 TMEM alloc, 3 idle, 4..7 drain/select
// DR = 2 adds warps 8..11 as a second drain group: each group scans half of every tile's columns into its own candidate
// sets, so every SM sub-partition has two drain warps to hide one another's insert latency (see sim_topk_kernel)
constexpr int sim_threads(int dr) { return 128 + 128 * dr; }
constexpr int SIM_A_CHUNK_BYTES = SIM_BLOCK_M * SIM_BLOCK_K * 2;
constexpr uint32_t kNoRow = 0xFFFFFFFFu;

// ---------------------------------------------------------------------------------------------
// K11: one warp per row.  search[r] = half( x / (||x|| + 1e-8) ), zero-padded to dpad columns.
// The search copy only has to rank candidates; exact scores come from the master copy later, so
// the norm here is an ordinary warp-shuffle reduction.  What the exact pass does need is a bound on
// how far the rounded row is from the unit row it stands for: rvmax <- max(rvmax, ||x*inv - half(x*inv)||_2)
// (float bits, non-negative, so an integer atomicMax orders them; see the certificate in K9).
// ---------------------------------------------------------------------------------------------
B200_DEVINL void publish_row_error(float err_sq, uint32_t* rvmax) {
  const uint32_t bits = __float_as_uint(sqrtf(err_sq) * 1.001f + 1e-9f);
  if (bits > __ldcg(rvmax)) atomicMax(rvmax, bits);
}

template <typename TIn>
__global__ void __launch_bounds__(256) ingest_rows_kernel(const TIn* __restrict__ in, int64_t n, int d, int dpad,
                                                          __half* __restrict__ search, int64_t row0,
                                                          TIn* __restrict__ master, uint32_t* __restrict__ rvmax) {
  const int64_t r = (int64_t)blockIdx.x * (blockDim.x / kWarp) + (threadIdx.x / kWarp);
  if (r >= n) return;
  const int lane = threadIdx.x & 31;
  const TIn* src = in + r * (int64_t)d;
  float ss = 0.f;
  for (int i = lane; i < d; i += kWarp) {
    float x = static_cast<float>(src[i]);
    ss = fmaf(x, x, ss);
  }
  ss = warp_sum(ss);
  const float inv = 1.0f / (sqrtf(ss) + 1e-8f);
  __half* dst = search + (row0 + r) * (int64_t)dpad;
  TIn* mdst = master ? master + (row0 + r) * (int64_t)d : nullptr;
  float es = 0.f;
  for (int i = lane; i < dpad; i += kWarp) {
    float x = (i < d) ? static_cast<float>(src[i]) : 0.f;
    const float u = x * inv;
    const __half h = __float2half_rn(u);
    const float e = u - __half2float(h);
    es = fmaf(e, e, es);
    dst[i] = h;
    if (mdst && i < d) mdst[i] = src[i];
  }
  es = warp_sum(es);
  if (lane == 0) publish_row_error(es, rvmax);
}

// Vectorised K11 for d % 8 == 0 and d == dpad (the 384/768-d production shapes): 16-byte loads
// and stores, one warp per row, rows of 384 fp32 = 3 x float4 per lane.
__global__ void __launch_bounds__(256) ingest_rows_f32_vec_kernel(const float* __restrict__ in, int64_t n, int d,
                                                                  __half* __restrict__ search, int64_t row0,
                                                                  float* __restrict__ master,
                                                                  uint32_t* __restrict__ rvmax) {
  const int64_t r = (int64_t)blockIdx.x * (blockDim.x / kWarp) + (threadIdx.x / kWarp);
  if (r >= n) return;
  const int lane = threadIdx.x & 31;
  const float4* src = reinterpret_cast<const float4*>(in + r * (int64_t)d);
  const int nv = d >> 2;
  float ss = 0.f;
  for (int i = lane; i < nv; i += kWarp) {
    float4 x = __ldg(src + i);
    ss = fmaf(x.x, x.x, ss); ss = fmaf(x.y, x.y, ss); ss = fmaf(x.z, x.z, ss); ss = fmaf(x.w, x.w, ss);
  }
  ss = warp_sum(ss);
  const float inv = 1.0f / (sqrtf(ss) + 1e-8f);
  uint2* dst = reinterpret_cast<uint2*>(search + (row0 + r) * (int64_t)d);
  float4* mdst = master ? reinterpret_cast<float4*>(master + (row0 + r) * (int64_t)d) : nullptr;
  float es = 0.f;
  for (int i = lane; i < nv; i += kWarp) {
    float4 x = __ldg(src + i);
    const float u0 = x.x * inv, u1 = x.y * inv, u2 = x.z * inv, u3 = x.w * inv;
    __half2 lo = __floats2half2_rn(u0, u1);
    __half2 hi = __floats2half2_rn(u2, u3);
    const float2 flo = __half22float2(lo), fhi = __half22float2(hi);
    const float e0 = u0 - flo.x, e1 = u1 - flo.y, e2 = u2 - fhi.x, e3 = u3 - fhi.y;
    es = fmaf(e0, e0, es); es = fmaf(e1, e1, es); es = fmaf(e2, e2, es); es = fmaf(e3, e3, es);
    uint2 o;
    o.x = *reinterpret_cast<uint32_t*>(&lo);
    o.y = *reinterpret_cast<uint32_t*>(&hi);
    dst[i] = o;
    if (mdst) mdst[i] = x;
  }
  es = warp_sum(es);
  if (lane == 0) publish_row_error(es, rvmax);
}

// ---------------------------------------------------------------------------------------------
// canonical fp32 arithmetic shared with oracle/search_oracle.c (bit-for-bit):
//   element i accumulates into partial (i & 31) by fused multiply-add, ascending i; partials are
//   folded 16, 8, 4, 2, 1.  One warp; every lane returns the same value.
// ---------------------------------------------------------------------------------------------
template <typename TA>
B200_DEVINL float canon_dot_self(const TA* __restrict__ x, int d, int lane) {
  float p = 0.f;
  for (int i = lane; i < d; i += kWarp) {
    const float v = static_cast<float>(x[i]);
    p = __fmaf_rn(v, v, p);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) p = __fadd_rn(p, __shfl_xor_sync(0xffffffffu, p, o));
  return p;
}

// score of raw row v against the already-normalised query qh (shared memory, fp32)
template <typename TV>
B200_DEVINL float canon_score(const float* __restrict__ qh, const TV* __restrict__ v, int d, int lane) {
  const float den = __fadd_rn(__fsqrt_rn(canon_dot_self(v, d, lane)), 1e-8f);
  float p = 0.f;
  for (int i = lane; i < d; i += kWarp) {
    const float vh = __fdiv_rn(static_cast<float>(v[i]), den);
    p = __fmaf_rn(qh[i], vh, p);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) p = __fadd_rn(p, __shfl_xor_sync(0xffffffffu, p, o));
  return p;
}

// ---------------------------------------------------------------------------------------------
// queries: raw (fp32 or bf16) -> half [nq_pad, dpad] tile operand (zero padded) and fp32 [nq, d]
// for the rescoring pass.  Queries are NOT normalised for the tensor-core pass: a positive
// per-query scale does not change that query's ranking.  They are scaled by 1/max|q| so that the
// half conversion neither overflows nor loses small-magnitude inputs to denormals.
//
// qmeta[r] = (inv_scale, r_q) for the certificate in K9:
//   inv_scale = 1 / (sc * (||q|| + 1e-8))     tensor-core score * inv_scale = score in cosine units
//   r_q       = || q_hat - q16 * inv_scale ||_2   distance between the query the exact pass uses
//               (q_hat = q / (||q|| + 1e-8)) and the one the tensor cores saw, in the same units
// ---------------------------------------------------------------------------------------------
template <typename TIn>
__global__ void __launch_bounds__(256) prep_queries_kernel(const TIn* __restrict__ in, int nq, int nq_pad, int d,
                                                           int dpad, __half* __restrict__ q16,
                                                           float* __restrict__ q_f32, float2* __restrict__ qmeta) {
  const int r = blockIdx.x * (blockDim.x / kWarp) + (threadIdx.x / kWarp);
  if (r >= nq_pad) return;
  const int lane = threadIdx.x & 31;
  __half* dst = q16 + (int64_t)r * dpad;
  if (r >= nq) {
    for (int i = lane; i < dpad; i += kWarp) dst[i] = __float2half_rn(0.f);
    return;
  }
  const TIn* src = in + (int64_t)r * d;
  float amax = 0.f;
  for (int i = lane; i < d; i += kWarp) amax = fmaxf(amax, fabsf(static_cast<float>(src[i])));
  amax = warp_max(amax);
  const float sc = (amax > 0.f && amax < INFINITY) ? (1.0f / amax) : 1.0f;
  const float den = __fadd_rn(__fsqrt_rn(canon_dot_self(src, d, lane)), 1e-8f);
  const float inv_scale = 1.0f / (sc * den);
  float es = 0.f;
  for (int i = lane; i < dpad; i += kWarp) {
    float x = (i < d) ? static_cast<float>(src[i]) : 0.f;
    const __half h = __float2half_rn(x * sc);
    dst[i] = h;
    if (i < d) {
      q_f32[(int64_t)r * d + i] = x;
      const float e = __fdiv_rn(x, den) - __half2float(h) * inv_scale;
      es = fmaf(e, e, es);
    }
  }
  es = warp_sum(es);
  if (qmeta && lane == 0) qmeta[r] = make_float2(inv_scale, sqrtf(es) * 1.001f + 1e-9f);
}

// ---------------------------------------------------------------------------------------------
// K8: similarity tiles with fused top-KP selection.
// ---------------------------------------------------------------------------------------------
struct SimParams {
  int a_stream;      // 1: the query tile does not fit in shared memory (dim > 384): its 64-wide chunks travel with
                     //    the corpus chunks through the ring (re-read from L2 per corpus tile) instead of staying resident
  int nq;                 // valid queries
  int kc;                 // 64-wide K chunks per row (dpad / 64), 1..6 with the query tile resident
  int nstage;             // corpus-tile pipeline stages
  uint32_t n_rows;        // valid corpus rows
  uint32_t tiles_per_split;
  int nsplit;
  uint64_t* cand;         // [nq][nsplit][KP] packed (score, row) keys, 0 = empty slot
  uint32_t* gbkt;         // [nq][KP] per-query, per-bucket best score seen by any CTA of the bucket
                          // (orderable bits, 0 = nothing yet), zeroed before the launch; see the drain
  int use_bound;          // 1: the bucket bound is on
  int bkt_per_cta;        // buckets a CTA feeds: tile t of split s reports into bucket (s * bkt_per_cta + t % bkt_per_cta) % KP,
                          // so that KP buckets exist (over disjoint rows) even when there are fewer than KP splits
  uint32_t* gthr;         // [nq] best "worst kept score" of any CTA whose set is full (orderable bits, 0 = none)
  uint32_t refresh_mask;  // the shared bounds are re-read every (refresh_mask + 1) tiles
  int use_gthr;
};

// One search hit as a 16-byte record: x = orderable(score) << 32 (0 = no hit), y = global id (-1 = no hit).
// This is what travels between GPUs: a shard's K9 writes records straight into the all-gather send buffer and
// K10 reads the gathered buffer, with nothing in between.
// Where a packed result goes: one buffer, or the same slot of a buffer on EVERY GPU of the group (peer-mapped
// pointers: plain stores over NVLink), which replaces the all-gather of the sharded search (see sharded.py).
constexpr int kMaxPeers = 8;
struct HitSink {
  ulonglong2* one;                 // nullptr: not a packed result (ids / scores arrays) unless n_peers > 0
  ulonglong2* peer[kMaxPeers];     // n_peers > 0: this shard's slot in every rank's receive buffer
  int n_peers;
};

B200_DEVINL void write_hit(int64_t* ids, float* scores, const HitSink& sink, size_t at, uint64_t key, int64_t id_base,
                           bool negate_score) {
  const bool hit = key != 0;
  const int64_t id = hit ? id_base + (int64_t)(~(uint32_t)(key & 0xFFFFFFFFu)) : -1;
  if (sink.one || sink.n_peers) {
    ulonglong2 r;
    r.x = hit ? (key & 0xFFFFFFFF00000000ull) : 0ull;
    r.y = (unsigned long long)id;
    if (sink.one) sink.one[at] = r;
#pragma unroll
    for (int i = 0; i < kMaxPeers; ++i)
      if (i < sink.n_peers) sink.peer[i][at] = r;
  } else {
    const float sc = orderable_f32((uint32_t)(key >> 32));
    ids[at] = id;
    scores[at] = hit ? (negate_score ? -sc : sc) : 0.f;
  }
}

// (score, row) -> one 64-bit key whose unsigned order is "better first": higher score, then lower
// row.  0 is reserved for "empty slot" (no real key is 0: orderable(-inf) is already 0x007FFFFF).
B200_DEVINL uint64_t pack_key(float s, uint32_t id) {
  return ((uint64_t)f32_orderable(s) << 32) | (uint64_t)(~id);
}

// v[j] for a run-time j without sending the array to local memory: a binary tree of selects (31 of them)
B200_DEVINL float pick32(const uint32_t (&v)[32], int j) {
  uint32_t a[16], b[8], c[4], d[2];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = (j & 16) ? v[16 + i] : v[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) b[i] = (j & 8) ? a[8 + i] : a[i];
#pragma unroll
  for (int i = 0; i < 4; ++i) c[i] = (j & 4) ? b[4 + i] : b[i];
#pragma unroll
  for (int i = 0; i < 2; ++i) d[i] = (j & 2) ? c[2 + i] : c[i];
  return __uint_as_float((j & 1) ? d[1] : d[0]);
}

// One query's candidate set: KP unsorted keys in shared memory (a thread-private row) plus the
// position of the current worst one.  A new candidate overwrites the worst, then the worst is
// found again -- no shifting, so an insert costs the same whatever its rank (a sorted list made
// inserts the bottleneck once every TMEM lane carries a live query).  Rows arrive in ascending order
// and the caller only inserts when score > worst score, so among equal scores the earlier row stays:
// the reference's stable-sort tie rule (inmemory.py:118).
// The worst key is the lowest score, then the highest row.  Scores decide it almost always, so only the keys'
// HIGH words are loaded (shared-space loads: through a generic pointer the compiler emitted generic LD / ST), their
// minimum comes from a tree of 32-bit min (one instruction per node; the 64-bit compare-and-select with index tracking
// was nine), and the low words are looked at only when two kept rows share the lowest score.  ~90 instructions per
// insert instead of 175 -- the drain warp is alone on its SM sub-partition, so instructions are what an insert costs.
// `keys_s` = shared-space byte address of the KP keys.
// Returns (float bits of the worst score, -inf while the set is not full) << 32 | worst position.
template <int KP>
__device__ __noinline__ uint64_t topk_replace_worst(uint32_t keys_s, int wpos, float s, uint32_t id) {
  asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(keys_s + (uint32_t)wpos * 8u), "r"(~id), "r"(f32_orderable(s)) : "memory");
  uint32_t h[KP], m[KP];
#pragma unroll
  for (int i = 0; i < KP; ++i) {
    h[i] = lds32(keys_s + (uint32_t)i * 8u + 4u);
    m[i] = h[i];
  }
#pragma unroll
  for (int stride = 1; stride < KP; stride <<= 1) {
#pragma unroll
    for (int i = 0; i + stride < KP; i += 2 * stride) m[i] = min(m[i], m[i + stride]);
  }
  const uint32_t whi = m[0];
  // positions that hold the minimum, as a bit mask (two words: KP goes up to 64), built from independent partial masks
  uint32_t eq4[4] = {0u, 0u, 0u, 0u};
#pragma unroll
  for (int i = 0; i < KP; ++i) eq4[(i >> 5) * 2 + (i & 1)] |= (h[i] == whi ? 1u : 0u) << (i & 31);
  const uint64_t eq = ((uint64_t)(eq4[2] | eq4[3]) << 32) | (eq4[0] | eq4[1]);
  int wp = __ffsll((long long)eq) - 1;
  if (whi != 0 && (eq & (eq - 1ull)) != 0) {    // several kept rows with this score (not empty slots: any of those will do)
    uint32_t wlo = 0xFFFFFFFFu;
    for (uint64_t e = eq; e; e &= e - 1ull) {
      const int i = __ffsll((long long)e) - 1;
      const uint32_t lo = lds32(keys_s + (uint32_t)i * 8u);
      if (lo <= wlo) {                           // low word = ~row: the highest row is the worst
        wlo = lo;
        wp = i;
      }
    }
  }
  const float thr = whi ? orderable_f32(whi) : -INFINITY;
  return ((uint64_t)__float_as_uint(thr) << 32) | (uint32_t)wp;
}

#ifdef B200_K8_TIMING
// Development aid (not compiled into the shipped library): per corpus tile of CTA 0, the clocks its first drain warp
// spends scanning the accumulators ([t], t < 48) and waiting for them ([64 + t]).
__device__ unsigned long long g_k8_timing[128];
#endif

// CG = 1: one CTA per (query tile, corpus slice).  CG = 2: a CTA pair (cluster of 2, cta_group::2)
// per (two query tiles, corpus slice): each CTA keeps its own 128 queries resident and loads only
// HALF of every corpus tile; the leader's MMA (M = 256) drives both SMs' tensor cores and each CTA
// drains the scores of its own queries from its own TMEM.  Half the corpus bytes per SM per flop
// means the same shared-memory ring hides twice the latency and the L2 -> SM path carries half the
// traffic -- the two things that kept the single-CTA kernel at ~60 % of the tensor roofline.
//
// DR = 2: two drain groups (warps 4..7 and 8..11).  Group g scans columns [g * BLOCK_N / 2, +BLOCK_N / 2) of every tile
// into its own KP-sets and hands them over as their own candidate list ([query][split][group][KP]: to K9 the launch has
// nsplit * DR splits).  One drain warp per SM sub-partition cannot hide the latency of its own inserts (~450 clocks
// each, dependent shared-memory round trips); with independent queries the 32 lanes of a warp pass their thresholds at
// different columns, so while sets warm up -- the first dozen tiles of every CTA -- the drain, not the tensor pipe, sets
// the pace.  Costs one more candidate-set area of shared memory (two ring stages of the CTA-pair kernel), so the host
// uses it on small shards only, where the warm-up is a large share of the launch.
template <int BLOCK_N, int KP, int CG, int DR = 1>
__global__ void __launch_bounds__(sim_threads(DR), 1)
sim_topk_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_c,
                const SimParams p) {
  static_assert(DR == 1 || DR == 2, "one or two drain groups");
  static_assert((BLOCK_N / 32 / DR) % 2 == 0, "every drain group walks an even number of 32-column chunks");
  constexpr uint32_t B_ROWS = BLOCK_N / CG;                     // corpus rows this CTA loads per tile
  constexpr uint32_t B_STAGE_BYTES = B_ROWS * SIM_BLOCK_K * 2;
  constexpr uint32_t IDESC = umma_idesc_f16_f32(SIM_BLOCK_M * CG, BLOCK_N);   // half x half -> fp32
  static_assert(CG == 1 || CG == 2, "CTA group is 1 or 2");
  constexpr uint32_t TMEM_COLS = 2 * BLOCK_N;  // two accumulator stages
  static_assert(TMEM_COLS == 256 || TMEM_COLS == 512, "TMEM allocation must be a power of two");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  // resident mode: [query tile: kc chunks][ring: nstage corpus chunks]; streamed mode: [ring: nstage x (query
  // chunk + corpus chunk)]
  const uint32_t a_in_stage = p.a_stream ? SIM_A_CHUNK_BYTES : 0u;
  const uint32_t stage_stride = a_in_stage + B_STAGE_BYTES;
  uint8_t* sA = smem;
  uint8_t* sB = p.a_stream ? smem : sA + (size_t)p.kc * SIM_A_CHUNK_BYTES;
  uint64_t* ckeys = reinterpret_cast<uint64_t*>(sB + (size_t)p.nstage * stage_stride);   // [DR][128][KP+1]
  uint64_t* bars = ckeys + DR * SIM_BLOCK_M * (KP + 1);
  uint64_t* full = bars;                       // [nstage]  TMA -> MMA
  uint64_t* empty = bars + p.nstage;           // [nstage]  MMA -> TMA
  uint64_t* acc_full = bars + 2 * p.nstage;    // [2]       MMA -> drain
  uint64_t* acc_empty = acc_full + 2;          // [2]       drain -> MMA
  uint64_t* a_full = acc_empty + 2;            // [1]       query tile landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(a_full + 1);

  const int warp = uniform_warp_idx();
  const int lane = threadIdx.x & 31;
  const int qtile = blockIdx.x;                     // CTA pair = query tiles (2i, 2i+1); CTAs sharing a corpus slice run together
  const int split = blockIdx.y;
  const int rank = (CG == 2) ? (int)cluster_ctarank() : 0;
  const bool leader = rank == 0;

  const uint32_t total_tiles = (p.n_rows + BLOCK_N - 1) / BLOCK_N;
  const uint32_t tile0 = (uint32_t)split * p.tiles_per_split;
  const uint32_t ntiles = tile0 < total_tiles ? min(p.tiles_per_split, total_tiles - tile0) : 0u;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_q);
    tma_prefetch_desc(&map_c);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < p.nstage; ++i) {
      mbar_init(&full[i], CG);      // one arrival per CTA of the group (the leader's copy is used)
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], 4 * CG * DR);  // one arrival per drain warp of the CTA group (leader's copy)
    }
    mbar_init(a_full, CG);
    fence_mbar_init();
  }
  if (warp == 2) {
    if (CG == 2) tmem_alloc_2sm(tmem_slot, TMEM_COLS);
    else tmem_alloc(tmem_slot, TMEM_COLS);
  }
  // candidate sets start empty
  for (int i = threadIdx.x; i < DR * SIM_BLOCK_M * (KP + 1); i += sim_threads(DR)) ckeys[i] = 0;
  tc_fence_before();
  if (CG == 2) cluster_sync_all();   // the peer's barriers exist before anyone signals them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer: the warp runs the loop converged, one elected lane issues (see uniform_warp_idx) =====
    if (ntiles > 0) {
      if (!p.a_stream && elect_one()) {
        if (CG == 2) {
          if (leader) mbar_arrive_expect_tx(a_full, (uint32_t)p.kc * SIM_A_CHUNK_BYTES * CG);
          else mbar_arrive_leader(a_full);
          for (int c = 0; c < p.kc; ++c)
            tma_load_2d_2sm(sA + (size_t)c * SIM_A_CHUNK_BYTES, &map_q, a_full, c * SIM_BLOCK_K, qtile * SIM_BLOCK_M);
        } else {
          mbar_arrive_expect_tx(a_full, (uint32_t)p.kc * SIM_A_CHUNK_BYTES);
          for (int c = 0; c < p.kc; ++c)
            tma_load_2d(sA + (size_t)c * SIM_A_CHUNK_BYTES, &map_q, a_full, c * SIM_BLOCK_K, qtile * SIM_BLOCK_M);
        }
      }
      __syncwarp();
      int s = 0;
      uint32_t ph = 0;
      for (uint32_t t = 0; t < ntiles; ++t) {
        const int32_t row = (int32_t)((tile0 + t) * BLOCK_N + rank * B_ROWS);
        for (int c = 0; c < p.kc; ++c) {
          mbar_wait(&empty[s], ph ^ 1);
          if (elect_one()) {
            uint8_t* stg = sB + (size_t)s * stage_stride;
            if (CG == 2) {
              if (leader) mbar_arrive_expect_tx(&full[s], stage_stride * CG);
              else mbar_arrive_leader(&full[s]);
              if (p.a_stream) tma_load_2d_2sm(stg, &map_q, &full[s], c * SIM_BLOCK_K, qtile * SIM_BLOCK_M);
              tma_load_2d_2sm(stg + a_in_stage, &map_c, &full[s], c * SIM_BLOCK_K, row);
            } else {
              mbar_arrive_expect_tx(&full[s], stage_stride);
              if (p.a_stream) tma_load_2d(stg, &map_q, &full[s], c * SIM_BLOCK_K, qtile * SIM_BLOCK_M);
              tma_load_2d(stg + a_in_stage, &map_c, &full[s], c * SIM_BLOCK_K, row);
            }
          }
          __syncwarp();
          if (++s == p.nstage) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the leader CTA's warp 1 runs the loop converged, one elected lane issues =====
    if (ntiles > 0 && leader) {
      if (!p.a_stream) mbar_wait(a_full, 0);
      tc_fence_after();
      const uint32_t a_addr = smem_u32(sA);
      const uint32_t b_addr = smem_u32(sB);
      int s = 0;
      uint32_t ph = 0;
      for (uint32_t t = 0; t < ntiles; ++t) {
        const uint32_t as = t & 1, aph = (t >> 1) & 1;
        mbar_wait(&acc_empty[as], aph ^ 1);
        tc_fence_after();
        const uint32_t tacc = tmem_base + as * BLOCK_N;
        for (int c = 0; c < p.kc; ++c) {
          mbar_wait(&full[s], ph);
          tc_fence_after();
          const uint64_t da = umma_smem_desc_sw128(p.a_stream ? b_addr + (uint32_t)s * stage_stride
                                                              : a_addr + (uint32_t)c * SIM_A_CHUNK_BYTES);
          const uint64_t db = umma_smem_desc_sw128(b_addr + (uint32_t)s * stage_stride + a_in_stage);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < SIM_BLOCK_K / 16; ++k) {
              // +32 bytes per K=16 step inside the 128-byte swizzle row: +2 in the >>4 address field
              if (CG == 2) umma_bf16_2sm(tacc, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), IDESC, (uint32_t)((c | k) != 0));
              else umma_bf16(tacc, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), IDESC, (uint32_t)((c | k) != 0));
            }
            if (CG == 2) umma_commit_2sm(&empty[s]); else umma_commit(&empty[s]);   // stage reusable once read
          }
          __syncwarp();
          if (++s == p.nstage) { s = 0; ph ^= 1; }
        }
        if (elect_one()) {
          if (CG == 2) umma_commit_2sm(&acc_full[as]); else umma_commit(&acc_full[as]);   // accumulator tile complete
        }
        __syncwarp();
      }
    }
  } else if (warp >= 4) {
    // ===== drain + select: thread owns one query (one TMEM lane) =====
    const int dg = (warp - 4) >> 2;            // drain group: which share of every tile's columns
    const int wq = (warp - 4) & 3;             // TMEM lane group this warp may read (= warp % 4)
    const int row = wq * 32 + lane;
    const int qg = qtile * SIM_BLOCK_M + row;
    uint64_t* my_keys = ckeys + (dg * SIM_BLOCK_M + row) * (KP + 1);      // odd stride in 8-byte words: conflict-free
    const uint32_t my_keys_s = smem_u32(my_keys);
    int wpos = 0;
    const bool live = qg < p.nq;
    // thr = max(score of this CTA's own worst kept candidate, just below the cross-CTA bound).
    //
    // The cross-CTA bound: the CTAs scanning one query tile are dealt into KP buckets (split % KP)
    // and every CTA publishes the best score it has inserted so far into its bucket's slot
    // (atomicMax).  Each non-empty slot is the score of a real row and the buckets cover disjoint
    // row ranges, so the MINIMUM over the KP slots has at least KP distinct rows at or above it: it
    // is a lower bound on the query's final KP-th best score.  Its quality is global -- the bucket
    // maxima are, roughly, the best rows of the whole corpus scanned so far -- whereas a CTA's own
    // KP-th best only reflects its own slice.  Without it half of all 32-column chunks of a fully
    // populated query tile reach the per-column insert path; with it a few percent do.  Until all
    // KP buckets have reported (the first waves of a multi-wave launch) the weaker bound gthr[q] --
    // the best KP-th-best of any single CTA, which alone holds KP rows at least that good -- stands
    // in.  Both bounds are
    // applied non-strictly (rows EQUAL to it still pass, hence "just below"): ties at the boundary
    // must all survive so that the lower row can win in the merge, whichever slice it sits in.
    float thr_own = live ? -INFINITY : INFINITY;     // padded query rows never insert
    float thr = thr_own;
    const bool bound_on = live && p.use_bound;
    uint32_t* my_bkt = p.gbkt + (size_t)(live ? qg : 0) * KP;
    for (uint32_t t = 0; t < ntiles; ++t) {
      const uint32_t as = t & 1, aph = (t >> 1) & 1;
      // the bucket this tile's rows report into, and the best score it is known to hold (a bucket's rows are disjoint
      // from every other bucket's; its best row either stays in some CTA's set or is pushed out by KP better ones)
      const int bkt = (int)(((uint32_t)split * (uint32_t)p.bkt_per_cta + t % (uint32_t)p.bkt_per_cta) % (uint32_t)KP);
      uint32_t bkt_best = bound_on ? __ldcg(my_bkt + bkt) : 0xFFFFFFFFu;
      if (live && (t & p.refresh_mask) == 0) {
        uint32_t gb = 0xFFFFFFFFu;
        if (!bound_on) gb = 0;
#pragma unroll
        for (int i = 0; i < KP / 4; ++i) {
          const uint4 b4 = __ldcg(reinterpret_cast<const uint4*>(my_bkt) + i);
          gb = min(min(gb, b4.x), min(b4.y, min(b4.z, b4.w)));
        }
        if (p.use_gthr) gb = max(gb, __ldcg(p.gthr + qg));   // bucket bound is 0 until every bucket has reported
        if (gb) {
          float below = orderable_f32(gb - 1);                  // previous float in the total order
          if (below == 0.f) below = __uint_as_float(0x80000001u);  // -0 compares equal to +0: step past it
          thr = fmaxf(thr, below);
        }
      }
#ifdef B200_K8_TIMING
      const long long kt0 = clock64();
#endif
      mbar_wait(&acc_full[as], aph);
      tc_fence_after();
#ifdef B200_K8_TIMING
      const long long kt1 = clock64();
#endif
      const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16) + as * BLOCK_N;
      const uint32_t col0 = (tile0 + t) * BLOCK_N;
      // Software-pipelined drain: the tcgen05.ld of chunk c+1 is in flight while chunk c is
      // scanned; the accumulator stage goes back to the MMA warp as soon as its last chunk is in
      // registers.
      constexpr int NCH = BLOCK_N / 32 / DR;     // chunks this drain group walks
      const int cbeg = dg * NCH;
      uint32_t v[2][32];
      tmem_ld32(taddr + cbeg * 32, v[0]);
#pragma unroll 1
      for (int c2 = 0; c2 < NCH; c2 += 2) {
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int c = cbeg + c2 + u;
          tmem_ld_wait_dep(v[u]);
          if (c2 + u + 1 < NCH) {
            tmem_ld32(taddr + (c + 1) * 32, v[u ^ 1]);
          } else {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              if (CG == 2) mbar_arrive_leader(&acc_empty[as]); else mbar_arrive(&acc_empty[as]);
            }
          }
          // maximum of the chunk by four independent chains (the one drain warp of an SM sub-partition has nobody to hide
          // a 31-deep FMNMX chain behind)
          float g[4];
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4) {
            const float a0 = fmaxf(__uint_as_float(v[u][q4 * 8 + 0]), __uint_as_float(v[u][q4 * 8 + 1]));
            const float a1 = fmaxf(__uint_as_float(v[u][q4 * 8 + 2]), __uint_as_float(v[u][q4 * 8 + 3]));
            const float a2 = fmaxf(__uint_as_float(v[u][q4 * 8 + 4]), __uint_as_float(v[u][q4 * 8 + 5]));
            const float a3 = fmaxf(__uint_as_float(v[u][q4 * 8 + 6]), __uint_as_float(v[u][q4 * 8 + 7]));
            g[q4] = fmaxf(fmaxf(a0, a1), fmaxf(a2, a3));
          }
          const float m = fmaxf(fmaxf(g[0], g[1]), fmaxf(g[2], g[3]));
          if (m > thr) {
            // Insert path.  Every lane makes the bit mask of ITS columns above the threshold and then pops them in
            // ascending order, so the warp runs max-over-lanes(popcount) rounds of ONE compact loop body.  (The first
            // version walked the 32 columns in straight-line code, one predicated insert per column: while a CTA's sets
            // warm up nearly every column has some lane above its threshold, and 32 call sites per chunk did not fit the
            // instruction cache -- the first tile of every CTA took 150 k clocks, ~600 per column.)
            const uint32_t base = col0 + c * 32;
            uint32_t mask = 0;
#pragma unroll
            for (int j = 0; j < 32; ++j) mask |= (__uint_as_float(v[u][j]) > thr ? 1u : 0u) << j;
            if (base + 32 > p.n_rows) mask &= base < p.n_rows ? (1u << (p.n_rows - base)) - 1u : 0u;   // rows past the end
            while (mask) {
              const int j = __ffs(mask) - 1;
              mask &= mask - 1;
              const float s = pick32(v[u], j);
              if (s > thr) {                      // the threshold has moved since the mask was made
                const uint64_t r = topk_replace_worst<KP>(my_keys_s, wpos, s, base + j);
                thr_own = __uint_as_float((uint32_t)(r >> 32));
                wpos = (int)(uint32_t)r;
                thr = fmaxf(thr, thr_own);
                if (p.use_gthr && thr_own > -INFINITY) atomicMax(p.gthr + qg, f32_orderable(thr_own));   // set is full
                const uint32_t os = f32_orderable(s);
                if (os > bkt_best) {
                  bkt_best = os;
                  atomicMax(my_bkt + bkt, os);
                }
              }
            }
          }
        }
      }
#ifdef B200_K8_TIMING
      if (blockIdx.x == 0 && warp == 4 && lane == 0 && t < 48) {
        g_k8_timing[t] = (unsigned long long)(clock64() - kt1);
        g_k8_timing[64 + t] = (unsigned long long)(kt1 - kt0);
      }
#endif
    }
    // this split's candidates for this query
    if (qg < p.nq) {
      const size_t o = (((size_t)qg * p.nsplit + split) * DR + dg) * KP;
#pragma unroll 4
      for (int j = 0; j < KP; ++j) p.cand[o + j] = my_keys[j];
    }
  }

  tc_fence_before();
  if (CG == 2) cluster_sync_all();   // no CTA leaves while its peer may still signal its barriers
  else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    if (CG == 2) tmem_dealloc_2sm(tmem_base, TMEM_COLS);
    else tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// descending bitonic sort of n (power of two) 64-bit keys in shared memory, whole CTA
B200_DEVINL void bitonic_sort_desc(uint64_t* keys, int n) {
  for (int k = 2; k <= n; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const uint64_t a = keys[i], b = keys[ixj];
          const bool desc = (i & k) == 0;
          if (desc ? (a < b) : (a > b)) {
            keys[i] = b;
            keys[ixj] = a;
          }
        }
      }
      __syncthreads();
    }
  }
}

// The same network for exactly 64 keys, run by ONE warp (32 compare-exchanges per stage = one per lane): the CTA-wide
// version spends eight warps' worth of loop and barrier instructions on it, and K9 -- 13 k warp instructions per
// query -- is bound by instruction issue.  Every thread of the CTA calls it; ends with __syncthreads.
B200_DEVINL void bitonic_sort_desc64_one_warp(uint64_t* keys) {
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
#pragma unroll
    for (int k = 2; k <= 64; k <<= 1) {
#pragma unroll
      for (int j = k >> 1; j > 0; j >>= 1) {
        const int i = ((lane & ~(j - 1)) << 1) | (lane & (j - 1));
        const int l = i | j;
        const uint64_t a = keys[i], b = keys[l];
        const bool desc = (i & k) == 0;
        if (desc ? (a < b) : (a > b)) {
          keys[i] = b;
          keys[l] = a;
        }
        __syncwarp();
      }
    }
  }
  __syncthreads();
}

// ---------------------------------------------------------------------------------------------
// K9: one CTA per query.
//   1. gather nsplit*KP (approx score, row) candidates as 64-bit keys (score desc, row asc),
//      bitonic-sort, keep the best KP
//   2. rescore those KP rows from the master copy with the reference arithmetic (canonical order)
//   3. order by (exact score desc, row asc) and emit k results: ids are id_base + row, -1 padded
// ---------------------------------------------------------------------------------------------
struct MergeParams {
  int nq, nsplit, kp, k, d;
  int ncand_pow2;
  uint32_t n_rows;
  int64_t id_base;
  const uint64_t* cand;     // [nq][nsplit][kp] packed keys from K8, 0 = empty
  const float* q_f32;       // [nq, d] raw queries
  const void* master;       // [n_rows, d] raw rows, fp32 or bf16
  int master_is_bf16;
  int64_t* out_ids;         // [nq, k]
  float* out_scores;        // [nq, k]
  HitSink out_packed;       // packed [nq, k] records (orderable(score) << 32, global id) instead of the two arrays
  // --- K8's shared score bounds as it left them (see SimParams): candidates below them cannot be among the KP best ---
  const uint32_t* gbkt;     // [nq][kp] bucket bests, nullptr or use_bound == 0: not used
  const uint32_t* gthr;     // [nq] best full-set worst, nullptr or use_gthr == 0: not used
  int use_bound, use_gthr;
  // --- candidate-sufficiency certificate (all nullptr: skipped) ---
  const float2* qmeta;      // [nq] (inv_scale, r_q) from prep_queries_kernel
  const uint32_t* rvmax;    // float bits: max over rows of || unit row - its half image ||_2
  float slop;               // bound on the fp32 accumulation / evaluation error of both scores
  uint32_t* fail_count;     // number of queries appended to fail_idx by this launch
  int32_t* fail_idx;        // [nq] queries whose result is NOT certified (re-run by exact_topk_kernel)
  unsigned long long* stats;  // [2] cumulative: queries checked, queries that failed the certificate
};

__global__ void __launch_bounds__(256) merge_rescore_kernel(const MergeParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint64_t* keys = reinterpret_cast<uint64_t*>(smem_raw);
  float* qh = reinterpret_cast<float*>(keys + p.ncand_pow2);  // [d]
  uint64_t* fin = reinterpret_cast<uint64_t*>(qh + ((p.d + 1) & ~1));  // [64]
  __shared__ float s_den;

  __shared__ int s_cnt;
  const int q = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int ncand = p.nsplit * p.kp;
  const size_t o = (size_t)q * ncand;
  // Gather with a filter.  Every CTA of K8 hands over a FULL set of KP keys, most of them left from its warm-up
  // (a set only ever replaces its worst entry), so nsplit * KP keys arrive -- up to 8192 -- of which a few dozen
  // matter.  K8's shared bounds, as they stand now that it has finished, are lower bounds on this query's KP-th
  // best score (>= KP distinct rows at or above them are in the sets): everything below goes, and the bitonic sort
  // runs over what is left, not over 8192 keys (0.15 ms -> 0.02 ms per launch at 128 queries).
  uint32_t lb = 0;
  if (p.use_bound && p.gbkt) {
    lb = 0xFFFFFFFFu;
    for (int i = 0; i < p.kp; ++i) lb = min(lb, __ldcg(p.gbkt + (size_t)q * p.kp + i));
  }
  if (p.use_gthr && p.gthr) lb = max(lb, __ldcg(p.gthr + q));
  if (lb) lb -= 1;                                     // K8 applies the bounds non-strictly (and steps past -0)
  if (threadIdx.x == 0) s_cnt = 0;
  __syncthreads();
  const int ncand_round = (ncand + blockDim.x - 1) / blockDim.x * blockDim.x;
  for (int i = threadIdx.x; i < ncand_round; i += blockDim.x) {
    const uint64_t key = (i < ncand) ? __ldcg(p.cand + o + i) : 0ull;
    const bool keep = key != 0 && (uint32_t)(key >> 32) >= lb;
    const uint32_t m = __ballot_sync(0xffffffffu, keep);
    int base = 0;
    if (lane == 0 && m) base = atomicAdd(&s_cnt, __popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (keep) keys[base + __popc(m & ((1u << lane) - 1u))] = key;
  }
  __syncthreads();
  const int nkept = s_cnt;
  int npow2 = 64;
  while (npow2 < nkept) npow2 <<= 1;
  for (int i = nkept + threadIdx.x; i < npow2; i += blockDim.x) keys[i] = 0;
  // q_hat = q / (||q|| + 1e-8)   (inmemory.py:108)
  const float* qraw = p.q_f32 + (size_t)q * p.d;
  if (warp == 0) {
    const float den = __fadd_rn(__fsqrt_rn(canon_dot_self(qraw, p.d, lane)), 1e-8f);
    if (lane == 0) s_den = den;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < p.d; i += blockDim.x) qh[i] = __fdiv_rn(qraw[i], s_den);
  for (int i = threadIdx.x; i < 64; i += blockDim.x) fin[i] = 0;
  __syncthreads();                 // keys[], qh[] and fin[] are written
  if (npow2 == 64) bitonic_sort_desc64_one_warp(keys);
  else bitonic_sort_desc(keys, npow2);  // both end with __syncthreads

  for (int c = warp; c < p.kp; c += nwarp) {
    const uint64_t key = keys[c];
    if (key == 0) continue;
    const uint32_t id = ~(uint32_t)(key & 0xFFFFFFFFu);
    float s;
    if (p.master_is_bf16)
      s = canon_score(qh, reinterpret_cast<const __nv_bfloat16*>(p.master) + (size_t)id * p.d, p.d, lane);
    else
      s = canon_score(qh, reinterpret_cast<const float*>(p.master) + (size_t)id * p.d, p.d, lane);
    if (lane == 0) fin[c] = pack_key(s, id);
  }
  __syncthreads();
  bitonic_sort_desc64_one_warp(fin);
  // Certificate.  K8 ranked rows by s16(v) = q16 . v16 (tensor cores, fp32 accumulate) and every row
  // outside the KP survivors has s16 <= A, the KP-th best.  In cosine units (x inv_scale), for any row v
  //   exact(v) - s16(v) = q_hat.(v_hat - v16) + (q_hat - q16').v16 + (evaluation error of both sides)
  //                    <= r_v + r_q (1 + r_v) + slop                      (Cauchy-Schwarz, |q_hat| <= 1)
  // with r_v <= rvmax kept by the ingest kernel and r_q computed by prep_queries_kernel from the values
  // actually used.  So no outsider can score above A + eps; if the k-th exact score among the survivors
  // is strictly above that, the survivors contain the exact top-k, ties included.  Otherwise the query
  // goes to exact_topk_kernel.  Fewer than KP candidates (corpus smaller than KP) or fewer than k
  // results means every row was rescored: certified.
  if (p.fail_count && threadIdx.x == 0) {
    const uint64_t last = keys[p.kp - 1];
    const uint64_t kth = fin[p.k - 1];
    bool ok = true;
    if (last != 0 && kth != 0) {
      const float2 qm = p.qmeta[q];
      const float rv = __uint_as_float(__ldcg(p.rvmax));
      const float a_norm = orderable_f32((uint32_t)(last >> 32)) * qm.x;
      const float eps = (rv + qm.y * (1.0f + rv) + p.slop) * 1.001f;
      const float e_k = orderable_f32((uint32_t)(kth >> 32));
      ok = (e_k - a_norm) > eps;   // NaN compares false: uncertified
    }
    atomicAdd(p.stats, 1ull);
    if (!ok) {
      atomicAdd(p.stats + 1, 1ull);
      p.fail_idx[atomicAdd(p.fail_count, 1u)] = q;
    }
  }
  for (int j = threadIdx.x; j < p.k; j += blockDim.x)
    write_hit(p.out_ids, p.out_scores, p.out_packed, (size_t)q * p.k + j, (j < 64) ? fin[j] : 0ull, p.id_base, false);
}

// ---------------------------------------------------------------------------------------------
// Exact top-k for the queries K9 could not certify (near-duplicate rows whose 16-bit images tie, or a
// k-th and KP-th neighbour closer than the rounding bound): the canonical fp32 score of EVERY row,
// i.e. exactly what the oracle computes, so the result is bit-identical by construction.  Launched
// after every K9 with a fixed grid; *fail_count == 0 (the common case) makes it return at once, so
// the query path needs no host round trip to decide.  One master-copy pass per failed query.
//
// Each warp scores rows blockIdx.x * nwarp + warp, + gridDim.x * nwarp, ... and keeps its best 64 keys
// (two per lane; a candidate replaces the warp's minimum, found again with a butterfly).  A CTA then
// sorts its warps' keys and writes its best 64; the last CTA to finish a query (atomic ticket) folds
// the gridDim.x lists 63 at a time and writes the k results.  METRIC_L2 = 1 scores by squared Euclidean
// distance instead (see exact_l2 below); keys order "better first" either way.
// ---------------------------------------------------------------------------------------------
constexpr int EXACT_THREADS = 512;
constexpr int EXACT_KEEP = 64;

struct ExactParams {
  const uint32_t* fail_count;   // nullptr: run all nq queries (query j = j)
  const int32_t* fail_idx;
  int nq;                       // upper bound on *fail_count / number of queries when fail_count is nullptr
  const float* q_f32;           // [nq_total, d] raw queries
  const void* master;
  int master_is_bf16;
  uint32_t n_rows;
  int d, k;
  int64_t id_base;
  uint32_t* fail2_count;        // SCREEN only: queries the screened pass could not settle (for the unscreened re-run)
  int32_t* fail2_idx;
  float screen_margin;
  int stage_keys;               // keys of shared memory for staging the per-CTA lists (>= 1024, normally gridDim.x * 64)
  uint64_t* part;               // [2][gridDim.x][EXACT_KEEP] scratch, alternating between consecutive queries
  uint32_t* ticket;             // [nq] zeroed before the launch; gridDim.x + 1 = that query's result is written
  int64_t* out_ids;
  float* out_scores;
  HitSink out_packed;           // see MergeParams
};

// FAISS IndexFlatL2 semantics (faiss_store.py:99-101,212-224): squared Euclidean distance, ascending; the
// store converts it to a score 1/(1+d).  Canonical order: element i -> partial i & 31 by fused multiply-add of
// the difference with itself, halving tree -- shared with oracle/search_oracle.c (l2 entry points).
template <typename TV>
B200_DEVINL float canon_l2(const float* __restrict__ q, const TV* __restrict__ v, int d, int lane) {
  float p = 0.f;
  for (int i = lane; i < d; i += kWarp) {
    const float t = __fsub_rn(q[i], static_cast<float>(v[i]));
    p = __fmaf_rn(t, t, p);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) p = __fadd_rn(p, __shfl_xor_sync(0xffffffffu, p, o));
  return p;
}

// canon_score / canon_l2 of one row, generic in d (the fallback of canon_rows8 below)
template <int METRIC_L2, typename TV>
B200_DEVINL float canon_row(const float* __restrict__ qh, const TV* __restrict__ v, int d, int lane) {
  return METRIC_L2 ? -canon_l2(qh, v, d, lane) : canon_score(qh, v, d, lane);
}

// eight consecutive elements as they come from memory (16 or 32 bytes), unpacked to fp32 when they are used
struct Raw8Bf16 { uint4 u; };
struct Raw8F32 { float4 a, b; };
B200_DEVINL void load8_raw(const __nv_bfloat16* p, Raw8Bf16& r) { r.u = __ldg(reinterpret_cast<const uint4*>(p)); }
B200_DEVINL void load8_raw(const float* p, Raw8F32& r) {
  r.a = __ldg(reinterpret_cast<const float4*>(p));
  r.b = __ldg(reinterpret_cast<const float4*>(p) + 1);
}
B200_DEVINL void unpack8(const Raw8Bf16& r, float (&x)[8]) {
  const uint32_t w[4] = {r.u.x, r.u.y, r.u.z, r.u.w};
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    x[2 * t] = __uint_as_float(w[t] << 16);
    x[2 * t + 1] = __uint_as_float(w[t] & 0xFFFF0000u);
  }
}
B200_DEVINL void unpack8(const Raw8F32& r, float (&x)[8]) {
  x[0] = r.a.x; x[1] = r.a.y; x[2] = r.a.z; x[3] = r.a.w; x[4] = r.b.x; x[5] = r.b.y; x[6] = r.b.z; x[7] = r.b.w;
}
template <typename TV> struct Raw8;
template <> struct Raw8<__nv_bfloat16> { using type = Raw8Bf16; static constexpr int kInFlight = 8; };
template <> struct Raw8<float> { using type = Raw8F32; static constexpr int kInFlight = 4; };

// The canonical 32 partial sums of a row, folded: lane 4r + m holds partials 8m .. 8m + 7 of row r.
// Halving tree of oracle/search_oracle.c: p[l] += p[l + 16] (lanes m <- m ^ 2), p[l] += p[l + 8] (m <- m ^ 1), then 4, 2, 1
// inside the lane.  All four lanes of a row end with the same value.
B200_DEVINL float fold_partials8(float (&a)[8]) {
#pragma unroll
  for (int t = 0; t < 8; ++t) a[t] = __fadd_rn(a[t], __shfl_xor_sync(0xffffffffu, a[t], 2));
#pragma unroll
  for (int t = 0; t < 8; ++t) a[t] = __fadd_rn(a[t], __shfl_xor_sync(0xffffffffu, a[t], 1));
#pragma unroll
  for (int t = 0; t < 4; ++t) a[t] = __fadd_rn(a[t], a[t + 4]);
  a[0] = __fadd_rn(a[0], a[2]);
  a[1] = __fadd_rn(a[1], a[3]);
  return __fadd_rn(a[0], a[1]);
}

// Cosine screen for the exhaustive kernel: one pass, one division per ROW -- q_hat . v / (|v| + 1e-8) in ordinary fp32
// (any summation order), within a few 1e-6 of the canonical score (which costs two passes and one IEEE division per
// ELEMENT: three quarters of the kernel's time).  Same lane layout as canon_rows8; all four lanes of a row return its value.
template <typename TV>
B200_DEVINL float screen_rows8(const float* __restrict__ qh, const TV* __restrict__ rows, int d, int lane, int n_valid) {
  using Raw = typename Raw8<TV>::type;
  constexpr int U = Raw8<TV>::kInFlight;
  const int r = lane >> 2, m = lane & 3;
  const bool on = r < n_valid;
  const TV* v = rows + (size_t)r * d;
  const int nch = d >> 3;
  float ss = 0.f, dot = 0.f;
  for (int c0 = m; c0 < nch; c0 += 4 * U) {
    Raw raw[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (on && c0 + 4 * u < nch) load8_raw(v + 8 * (c0 + 4 * u), raw[u]);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int c = c0 + 4 * u;
      if (on && c < nch) {
        float x[8];
        unpack8(raw[u], x);
        const float4 q0 = *reinterpret_cast<const float4*>(qh + 8 * c);
        const float4 q1 = *reinterpret_cast<const float4*>(qh + 8 * c + 4);
        const float qv[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          ss = fmaf(x[t], x[t], ss);
          dot = fmaf(qv[t], x[t], dot);
        }
      }
    }
  }
  ss += __shfl_xor_sync(0xffffffffu, ss, 1);
  dot += __shfl_xor_sync(0xffffffffu, dot, 1);
  ss += __shfl_xor_sync(0xffffffffu, ss, 2);
  dot += __shfl_xor_sync(0xffffffffu, dot, 2);
  return on ? dot / (sqrtf(ss) + 1e-8f) : -INFINITY;
}

// Eight consecutive rows per warp with 16-byte loads, bit-identical to canon_score / canon_l2 per row.
// Lane 4r + m works on row r and on the 8-element chunks c = 4k + m (k = 0, 1, ...): chunk c holds elements
// 8c .. 8c + 7, which the canonical order sends to partials 8m .. 8m + 7 in ascending k -- so each lane keeps eight
// of its row's 32 partial sums and adds to them in exactly the canonical sequence, while its loads (one per k, all
// independent) are 16 bytes wide and four lanes cover 64 contiguous bytes.  One row per warp with 2-byte loads kept
// ~1 KB in flight per warp and ran at a fifth of the HBM rate.  Needs d % 8 == 0; rows >= n_valid are skipped.
template <int METRIC_L2, typename TV>
B200_DEVINL float canon_rows8(const float* __restrict__ qh, const TV* __restrict__ rows, int d, int lane, int n_valid) {
  using Raw = typename Raw8<TV>::type;
  constexpr int U = Raw8<TV>::kInFlight;        // loads issued before the first one is consumed (32 registers' worth)
  const int r = lane >> 2, m = lane & 3;
  const bool on = r < n_valid;
  const TV* v = rows + (size_t)r * d;
  const int nch = d >> 3;                       // chunks per row
  float a[8], den = 1.f;
  if (!METRIC_L2) {
#pragma unroll
    for (int t = 0; t < 8; ++t) a[t] = 0.f;
    for (int c0 = m; c0 < nch; c0 += 4 * U) {
      Raw raw[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (on && c0 + 4 * u < nch) load8_raw(v + 8 * (c0 + 4 * u), raw[u]);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (on && c0 + 4 * u < nch) {
          float x[8];
          unpack8(raw[u], x);
#pragma unroll
          for (int t = 0; t < 8; ++t) a[t] = __fmaf_rn(x[t], x[t], a[t]);
        }
      }
    }
    den = __fadd_rn(__fsqrt_rn(fold_partials8(a)), 1e-8f);
  }
#pragma unroll
  for (int t = 0; t < 8; ++t) a[t] = 0.f;
  for (int c0 = m; c0 < nch; c0 += 4 * U) {
    Raw raw[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (on && c0 + 4 * u < nch) load8_raw(v + 8 * (c0 + 4 * u), raw[u]);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int c = c0 + 4 * u;
      if (on && c < nch) {
        float x[8];
        unpack8(raw[u], x);
        const float4 q0 = *reinterpret_cast<const float4*>(qh + 8 * c);
        const float4 q1 = *reinterpret_cast<const float4*>(qh + 8 * c + 4);
        const float qv[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          if (METRIC_L2) {
            const float df = __fsub_rn(qv[t], x[t]);
            a[t] = __fmaf_rn(df, df, a[t]);
          } else {
            a[t] = __fmaf_rn(qv[t], __fdiv_rn(x[t], den), a[t]);
          }
        }
      }
    }
  }
  const float s = fold_partials8(a);
  return METRIC_L2 ? -s : s;
}

B200_DEVINL uint64_t shfl_xor_u64(uint64_t v, int o) {
  const uint32_t lo = __shfl_xor_sync(0xffffffffu, (uint32_t)v, o);
  const uint32_t hi = __shfl_xor_sync(0xffffffffu, (uint32_t)(v >> 32), o);
  return ((uint64_t)hi << 32) | lo;
}

// max of a 64-bit key over the warp: the high words first, then the low words of the lanes that hold the maximum
B200_DEVINL uint64_t warp_max_u64(uint64_t v) {
  const uint32_t hi = (uint32_t)(v >> 32);
  const uint32_t mhi = __reduce_max_sync(0xffffffffu, hi);
  const uint32_t mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? (uint32_t)v : 0u);
  return ((uint64_t)mhi << 32) | mlo;
}

// One warp merges up to 32 sorted lists (better first, 0 = end), one per lane, into out[0 .. kk).  Keys are unique
// across the lists, so exactly one lane owns each winner and steps its cursor.  `len` = 0: this lane has no list.
__device__ __noinline__ void warp_merge_heads(const uint64_t* lst, int len, int kk, uint64_t* out, int lane) {
  int pos = 0;
  uint64_t head = len > 0 ? lst[0] : 0ull;
  for (int t = 0; t < kk; ++t) {
    const uint64_t wbest = warp_max_u64(head);
    if (lane == 0) out[t] = wbest;
    if (wbest != 0 && head == wbest) {
      ++pos;
      head = pos < len ? lst[pos] : 0ull;
    }
  }
}

// SCREEN (cosine only, d % 8 == 0): rank every row by the screen value instead, then compute the canonical score of the
// best 64 only and check that it settles the top-k: the 64th screen value must lie more than `screen_margin` below the
// k-th canonical score (no row outside the 64 can then reach it).  Queries for which it does not -- more than 64 - k
// rows within the margin of the k-th, i.e. a large cluster of near-duplicates -- are appended to a second list and
// re-run by the unscreened kernel.
template <int METRIC_L2, int SCREEN>
__global__ void __launch_bounds__(EXACT_THREADS, 1) exact_topk_kernel(const ExactParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint64_t* mkeys = reinterpret_cast<uint64_t*>(smem_raw);                  // [p.stage_keys]
  float* qh = reinterpret_cast<float*>(mkeys + p.stage_keys);               // [d]
  __shared__ float s_den;
  __shared__ uint32_t s_last;
  __shared__ uint64_t s_carry[2][EXACT_KEEP];
  __shared__ uint64_t s_wout[16][EXACT_KEEP];   // EXACT_THREADS / 32 warps
  const int nfail = p.fail_count ? (int)min(__ldcg(p.fail_count), (uint32_t)p.nq) : p.nq;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int j = 0; j < nfail; ++j) {
    const int q = p.fail_count ? p.fail_idx[j] : j;
    const float* qraw = p.q_f32 + (size_t)q * p.d;
    if (METRIC_L2) {
      for (int i = threadIdx.x; i < p.d; i += blockDim.x) qh[i] = qraw[i];
    } else {
      if (warp == 0) {
        const float den = __fadd_rn(__fsqrt_rn(canon_dot_self(qraw, p.d, lane)), 1e-8f);
        if (lane == 0) s_den = den;
      }
      __syncthreads();
      for (int i = threadIdx.x; i < p.d; i += blockDim.x) qh[i] = __fdiv_rn(qraw[i], s_den);
    }
    __syncthreads();
    // --- this warp's rows, eight at a time ---
    uint64_t k0 = 0, k1 = 0, wmin = 0;       // two kept keys per lane; wmin = the smallest of the 64
    const bool vec = (p.d & 7) == 0 && (reinterpret_cast<uintptr_t>(p.master) & 15) == 0;
    const uint32_t n_groups = (p.n_rows + 7) / 8;
    for (uint32_t grp = blockIdx.x * nwarp + warp; grp < n_groups; grp += gridDim.x * nwarp) {
      const uint32_t row0 = grp * 8;
      const int n_valid = (int)min(8u, p.n_rows - row0);
      float mine = 0.f;                       // lanes 4r .. 4r + 3: the score of row row0 + r
      if (SCREEN) {
        mine = p.master_is_bf16
                   ? screen_rows8(qh, reinterpret_cast<const __nv_bfloat16*>(p.master) + (size_t)row0 * p.d, p.d, lane, n_valid)
                   : screen_rows8(qh, reinterpret_cast<const float*>(p.master) + (size_t)row0 * p.d, p.d, lane, n_valid);
      } else if (vec) {
        mine = p.master_is_bf16
                   ? canon_rows8<METRIC_L2>(qh, reinterpret_cast<const __nv_bfloat16*>(p.master) + (size_t)row0 * p.d, p.d, lane, n_valid)
                   : canon_rows8<METRIC_L2>(qh, reinterpret_cast<const float*>(p.master) + (size_t)row0 * p.d, p.d, lane, n_valid);
      } else {
        for (int j = 0; j < n_valid; ++j) {
          const float sj = p.master_is_bf16
                               ? canon_row<METRIC_L2>(qh, reinterpret_cast<const __nv_bfloat16*>(p.master) + (size_t)(row0 + j) * p.d, p.d, lane)
                               : canon_row<METRIC_L2>(qh, reinterpret_cast<const float*>(p.master) + (size_t)(row0 + j) * p.d, p.d, lane);
          if ((lane >> 2) == j) mine = sj;
        }
      }
      for (int j = 0; j < n_valid; ++j) {
        const uint64_t key = pack_key(__shfl_sync(0xffffffffu, mine, 4 * j), row0 + j);
        if (key > wmin) {
          // replace one occurrence of the minimum (empty slots are 0 = the minimum until the set is full)
          const uint32_t has0 = __ballot_sync(0xffffffffu, k0 == wmin);
          const uint32_t has1 = __ballot_sync(0xffffffffu, k1 == wmin);
          if (has0) { if (lane == __ffs(has0) - 1) k0 = key; }
          else if (lane == __ffs(has1) - 1) k1 = key;
          uint64_t mn = k0 < k1 ? k0 : k1;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const uint64_t other = shfl_xor_u64(mn, o);
            mn = other < mn ? other : mn;
          }
          wmin = mn;
        }
      }
    }
    // --- CTA: best 64 of the warps' keys ---
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) mkeys[i] = 0;
    __syncthreads();
    mkeys[warp * 64 + lane] = k0;
    mkeys[warp * 64 + 32 + lane] = k1;
    __syncthreads();
    bitonic_sort_desc(mkeys, 1024);            // nwarp * 64 <= 1024
    // the list buffer of query j was last used by query j - 2: wait until that one has been folded
    // (the CTA doing it already holds its ticket, i.e. is running, so this cannot deadlock)
    if (j >= 2 && threadIdx.x == 0) {
      const volatile uint32_t* t2 = p.ticket + (j - 2);
      while (*t2 != gridDim.x + 1) __nanosleep(200);
      __threadfence();
    }
    __syncthreads();
    uint64_t* part = p.part + (size_t)(j & 1) * gridDim.x * EXACT_KEEP;
    if (threadIdx.x < EXACT_KEEP) part[(size_t)blockIdx.x * EXACT_KEEP + threadIdx.x] = mkeys[threadIdx.x];
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(p.ticket + j, 1u) == gridDim.x - 1) ? 1u : 0u;
    __syncthreads();
    if (s_last) {
      __threadfence();
      // Fold the per-CTA lists.  Each is sorted (better first) and only its first kk entries can reach the result.
      // They are staged in shared memory (normally all at once); every warp then merges 32 of them by their heads,
      // one list per lane (a round is a warp arg-max and one cursor step, ~100 clocks), and warp 0 merges the warps'
      // outputs with the carry of the previous round the same way.  (Sorting the gridDim.x * 64 keys with the bitonic
      // network cost 0.2 ms per query -- more than streaming a million rows.)
      const int kk = SCREEN ? EXACT_KEEP : min(p.k, EXACT_KEEP);
      const int per_round = min(p.stage_keys / kk, 32 * nwarp);
      int cur = 0;
      for (int i = threadIdx.x; i < kk; i += blockDim.x) s_carry[0][i] = 0;
      for (uint32_t c0 = 0; c0 < gridDim.x; c0 += per_round) {
        const int nl = (int)min((uint32_t)per_round, gridDim.x - c0);
        for (int i = threadIdx.x; i < nl * kk; i += blockDim.x) {
          const int l = i / kk, e = i - l * kk;
          mkeys[i] = __ldcg(part + (size_t)(c0 + l) * EXACT_KEEP + e);
        }
        __syncthreads();
        const int nw = (nl + 31) >> 5;
        if (warp < nw) {
          const int l = warp * 32 + lane;
          warp_merge_heads(mkeys + (size_t)l * kk, l < nl ? kk : 0, kk, s_wout[warp], lane);
        }
        __syncthreads();
        if (warp == 0)
          warp_merge_heads(lane == 0 ? s_carry[cur] : s_wout[(lane - 1) & 15], lane <= nw ? kk : 0, kk, s_carry[cur ^ 1], lane);
        cur ^= 1;
        __syncthreads();
      }
      for (int i = threadIdx.x; i < kk; i += blockDim.x) mkeys[i] = s_carry[cur][i];
      __syncthreads();
      bool settled = true;
      if (SCREEN) {
        // canonical scores of the 64 best rows by screen value; mkeys[0 .. 64) holds them in screen order
        __shared__ uint64_t s_exact[EXACT_KEEP];
        for (int c = warp; c < EXACT_KEEP; c += nwarp) {
          const uint64_t key = mkeys[c];
          uint64_t ek = 0;
          if (key) {
            const uint32_t id = ~(uint32_t)(key & 0xFFFFFFFFu);
            const float sc = p.master_is_bf16
                                 ? canon_score(qh, reinterpret_cast<const __nv_bfloat16*>(p.master) + (size_t)id * p.d, p.d, lane)
                                 : canon_score(qh, reinterpret_cast<const float*>(p.master) + (size_t)id * p.d, p.d, lane);
            ek = pack_key(sc, id);
          }
          if (lane == 0) s_exact[c] = ek;
        }
        __syncthreads();
        const uint64_t last_screen = mkeys[EXACT_KEEP - 1];        // 0: fewer than 64 rows exist, all of them are in
        bitonic_sort_desc(s_exact, EXACT_KEEP);
        const uint64_t kth = s_exact[min(p.k, EXACT_KEEP) - 1];
        if (last_screen != 0 && kth != 0) {
          const float e_k = orderable_f32((uint32_t)(kth >> 32));
          const float c_last = orderable_f32((uint32_t)(last_screen >> 32));
          settled = (e_k - c_last) > p.screen_margin;
        }
        for (int i = threadIdx.x; i < EXACT_KEEP; i += blockDim.x) mkeys[i] = s_exact[i];
        __syncthreads();
        if (!settled && threadIdx.x == 0) p.fail2_idx[atomicAdd(p.fail2_count, 1u)] = q;
      }
      if (settled) {
        for (int t = threadIdx.x; t < p.k; t += blockDim.x)
          write_hit(p.out_ids, p.out_scores, p.out_packed, (size_t)q * p.k + t, t < min(p.k, EXACT_KEEP) ? mkeys[t] : 0ull,
                    p.id_base, METRIC_L2 != 0);
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(p.ticket + j, 1u);           // gridDim.x + 1: this query's list buffer is free again
      }
    }
    __syncthreads();   // qh / mkeys are reused by the next query
  }
}

// ---------------------------------------------------------------------------------------------
// Exhaustive path for top_k > 48 (the fused select keeps at most 64 candidates per query and corpus
// slice).  The reference slices its sorted list at any top_k (inmemory.py:119) and one of its callers
// asks for 50 by default (orchestrator/app/context_sources.py:66), so large k must work -- it need not
// be fast: exact canonical fp32 score of EVERY row against one query (the arithmetic of the rescoring
// pass, so the scores are bit-identical to the oracle's), packed into the usual 64-bit keys, then
// repeated "sort 4096 keys in shared memory, keep the best 1024" passes until one list is left.
// HBM-bound on the master copy: ~0.3 ms per query per million 384-d fp32 rows.
// ---------------------------------------------------------------------------------------------
constexpr int EXH_CHUNK = 4096;      // keys sorted per CTA
constexpr int EXH_KEEP = 1024;       // survivors per CTA = the largest supported top_k

template <typename TV>
__global__ void __launch_bounds__(256) exh_score_kernel(const float* __restrict__ q_raw, const TV* __restrict__ master,
                                                        uint32_t n_rows, int d, uint64_t* __restrict__ keys) {
  extern __shared__ float exh_qh[];            // [d] normalised query
  __shared__ float s_den;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  if (warp == 0) {
    const float den = __fadd_rn(__fsqrt_rn(canon_dot_self(q_raw, d, lane)), 1e-8f);
    if (lane == 0) s_den = den;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < d; i += blockDim.x) exh_qh[i] = __fdiv_rn(q_raw[i], s_den);
  __syncthreads();
  for (uint32_t row = blockIdx.x * nwarp + warp; row < n_rows; row += gridDim.x * nwarp) {
    const float s = canon_score(exh_qh, master + (size_t)row * d, d, lane);
    if (lane == 0) keys[row] = pack_key(s, row);
  }
}

// keys_in[n_in] -> keys_out[gridDim.x * EXH_KEEP]: CTA b sorts keys [b * EXH_CHUNK, +EXH_CHUNK) (0 = none)
// in descending order and keeps the first EXH_KEEP.
__global__ void __launch_bounds__(512) exh_reduce_kernel(const uint64_t* __restrict__ keys_in, uint64_t n_in,
                                                         uint64_t* __restrict__ keys_out) {
  extern __shared__ uint64_t exh_keys[];       // [EXH_CHUNK]
  const uint64_t base = (uint64_t)blockIdx.x * EXH_CHUNK;
  for (int i = threadIdx.x; i < EXH_CHUNK; i += blockDim.x) exh_keys[i] = base + i < n_in ? keys_in[base + i] : 0ull;
  __syncthreads();
  bitonic_sort_desc(exh_keys, EXH_CHUNK);
  for (int i = threadIdx.x; i < EXH_KEEP; i += blockDim.x) keys_out[(uint64_t)blockIdx.x * EXH_KEEP + i] = exh_keys[i];
}

__global__ void exh_query_f32_kernel(const __nv_bfloat16* __restrict__ q, int d, float* __restrict__ out) {
  for (int i = threadIdx.x; i < d; i += blockDim.x) out[i] = __bfloat162float(q[i]);
}

// the single surviving sorted list -> ids / scores of one query
__global__ void exh_write_kernel(const uint64_t* __restrict__ keys, int n_keys, int k, int64_t id_base,
                                 int64_t* __restrict__ ids, float* __restrict__ scores, const HitSink packed, size_t at0) {
  for (int j = threadIdx.x; j < k; j += blockDim.x)
    write_hit(ids, scores, packed, at0 + (size_t)j, j < n_keys ? keys[j] : 0ull, id_base, false);
}

// "no rows": every slot of an [n] result is the empty hit
__global__ void fill_no_hits_kernel(int64_t* __restrict__ ids, float* __restrict__ scores, const HitSink packed,
                                    size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    write_hit(ids, scores, packed, i, 0ull, 0, false);
}

// ---------------------------------------------------------------------------------------------
// K10: merge nlists sorted (score, id) lists of k_in per query into the best k_out, global ids,
// (score desc, id asc).  Used after the all-gather of per-GPU results: lists arrive as
// [nlists][nq][k_in].  One warp per query; nlists*k_in <= 1024.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) merge_lists_kernel(const int64_t* __restrict__ ids_in,
                                                          const float* __restrict__ sc_in, int nq, int nlists,
                                                          int k_in, int k_out, int64_t* __restrict__ ids_out,
                                                          float* __restrict__ sc_out) {
  const int q = blockIdx.x * (blockDim.x / kWarp) + (threadIdx.x / kWarp);
  if (q >= nq) return;
  const int lane = threadIdx.x & 31;
  const int total = nlists * k_in;
  // selection by repeated warp arg-max over (score desc, id asc); total is small (<= 1024)
  float prev_s = INFINITY;
  int64_t prev_id = -1;
  for (int j = 0; j < k_out; ++j) {
    float best_s = -INFINITY;
    int64_t best_id = INT64_MAX;
    for (int i = lane; i < total; i += kWarp) {
      const int l = i / k_in, e = i - l * k_in;
      const size_t at = ((size_t)l * nq + q) * k_in + e;
      const int64_t id = ids_in[at];
      if (id < 0) continue;
      const float s = sc_in[at];
      // strictly after the previously emitted (prev_s, prev_id) in the total order
      const bool after = (s < prev_s) || (s == prev_s && id > prev_id);
      if (!after) continue;
      if (s > best_s || (s == best_s && id < best_id)) { best_s = s; best_id = id; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float os = __shfl_xor_sync(0xffffffffu, best_s, o);
      const int64_t oi = __shfl_xor_sync(0xffffffffu, best_id, o);
      if (os > best_s || (os == best_s && oi < best_id)) { best_s = os; best_id = oi; }
    }
    if (best_id == INT64_MAX) {
      if (lane == 0) { ids_out[(size_t)q * k_out + j] = -1; sc_out[(size_t)q * k_out + j] = 0.f; }
      prev_s = -INFINITY; prev_id = INT64_MAX;
    } else {
      if (lane == 0) { ids_out[(size_t)q * k_out + j] = best_id; sc_out[(size_t)q * k_out + j] = best_s; }
      prev_s = best_s; prev_id = best_id;
    }
  }
}

// K10 over packed records (see write_hit): lists arrive as [nlists][nq_per_list][k_in] -- exactly the buffer an
// all-gather of every shard's K9 output produces -- and this GPU merges queries [q_first, q_first + nq).
// Peer exchange, producer side: after the kernels that stored this shard's records into every rank's buffer
// (same stream, so their stores are complete), publish "search number `epoch` of rank `rank` has arrived" to every rank.
__global__ void signal_peers_kernel(const HitSink flags, int rank, uint32_t epoch) {
  const int i = threadIdx.x;
  if (i < flags.n_peers) {
    __threadfence_system();
    uint32_t* f = reinterpret_cast<uint32_t*>(flags.peer[i]) + rank;
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f), "r"(epoch) : "memory");
  }
}

// wait_flags != nullptr: consumer side of the peer exchange -- before reading, wait until every rank has published
// `epoch` (flags only grow; a rank may already be one search ahead, which uses the other half of the buffer).
__global__ void __launch_bounds__(128) merge_packed_kernel(const ulonglong2* __restrict__ in, int nlists, int nq_per_list,
                                                           int k_in, int q_first, int nq, int k_out,
                                                           int64_t* __restrict__ ids_out, float* __restrict__ sc_out,
                                                           const uint32_t* __restrict__ wait_flags, uint32_t epoch) {
  const int q = blockIdx.x * (blockDim.x / kWarp) + (threadIdx.x / kWarp);
  const int lane = threadIdx.x & 31;
  if (wait_flags) {
    if (lane < nlists) {
      uint32_t v;
      do {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(wait_flags + lane) : "memory");
      } while ((int32_t)(v - epoch) < 0);
    }
    __syncwarp();
  }
  if (q >= nq) return;
  const int total = nlists * k_in;
  // (x desc, id asc) as one comparison: x carries the orderable score in its high half, ids are non-negative
  unsigned long long prev_x = ~0ull;
  long long prev_id = -1;
  for (int j = 0; j < k_out; ++j) {
    unsigned long long best_x = 0ull;
    long long best_id = INT64_MAX;
    for (int i = lane; i < total; i += kWarp) {
      const int l = i / k_in, e = i - l * k_in;
      const ulonglong2 r = __ldcg(in + ((size_t)l * nq_per_list + q_first + q) * k_in + e);   // not the read-only path: peers wrote it
      const long long id = (long long)r.y;
      if (id < 0) continue;
      const bool after = (r.x < prev_x) || (r.x == prev_x && id > prev_id);
      if (!after) continue;
      if (r.x > best_x || (r.x == best_x && id < best_id)) { best_x = r.x; best_id = id; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long ox = shfl_xor_u64(best_x, o);
      const long long oi = (long long)shfl_xor_u64((uint64_t)best_id, o);
      if (ox > best_x || (ox == best_x && oi < best_id)) { best_x = ox; best_id = oi; }
    }
    const size_t at = (size_t)q * k_out + j;
    if (best_id == INT64_MAX) {
      if (lane == 0) { ids_out[at] = -1; sc_out[at] = 0.f; }
      prev_x = 0ull; prev_id = INT64_MAX;
    } else {
      if (lane == 0) { ids_out[at] = best_id; sc_out[at] = orderable_f32((uint32_t)(best_x >> 32)); }
      prev_x = best_x; prev_id = best_id;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// row gather (for SearchResult.vector / get()) and stable compaction (delete)
// ---------------------------------------------------------------------------------------------
template <typename TIn, typename TOut>
__global__ void gather_rows_kernel(const TIn* __restrict__ src, const int64_t* __restrict__ rows, int64_t n, int d,
                                   TOut* __restrict__ dst) {
  const int64_t r = blockIdx.x;
  if (r >= n) return;
  const TIn* s = src + rows[r] * (int64_t)d;
  TOut* o = dst + r * (int64_t)d;
  for (int i = threadIdx.x; i < d; i += blockDim.x) o[i] = static_cast<TOut>(static_cast<float>(s[i]));
}

}  // namespace b200
