// attention_tc64.cuh -- K3 on tcgen05 for 64-wide heads and sequences of whole 128-key blocks (the bert-base /
// mpnet-base-class shape: 12 heads x 64, seq 384): softmax(QK^T/sqrt(dh) + key-padding mask) V.
//
// Same arithmetic as attention_tc.cuh (read that file first): scores with the softmax reference already
// subtracted by the tensor core ([Q | -m] [K | 1]^T), probabilities 2^(s') written back over the scores as bf16 and
// used as the A operand straight from TMEM, row sums from a block of ones appended to V.  What differs is the
// shape of the work.  A 128 x 384 fp32 score tile is 384 TMEM columns, so two of them (two softmax groups) do not
// fit; and a (sequence, head) has 48 KB of keys and 48 KB of values, so whole K / V per pipeline stage do not fit
// either.  Hence:
//
//   key blocks   a query tile walks its keys in blocks of 128: S_b = Q K_b^T (128 TMEM columns), P_b = 2^(S_b),
//                O += P_b V_b.  The softmax reference is the Cauchy-Schwarz bound |q_r| max_j |k_j| over ALL keys
//                of the (sequence, head), fixed before the first block, so the blocks simply accumulate: no
//                running maximum, no rescaling of O.  max_j |k_j|^2 comes from a small kernel that runs first
//                (att_norms_kernel); it also decides, for the whole launch, whether the bound is usable.
//   K / V ring   K_b and V_b (16 KB each) travel through a ring of three 32 KB slots in the order the MMA
//                issuer consumes them; the two groups work on different query tiles in lock step.
//   TMEM         per group: 128 score columns + 80 output columns (64 dims, then the row sum 16 times).
//   exact mode   when any (sequence, head) has 2 |q| |k| above ATT_BOUND_GAP (decided by att_norms_kernel, one flag
//                for the launch, so that every warp role follows the same script), each tile first walks its key
//                blocks once for the exact row maximum (scores only, K only), then again for the probabilities.
//
// Warp roles: 0-7 softmax group 0, 8-15 group 1 (within a group warp w: TMEM lane group w & 3, key half w >> 2 of
// every block), 16 TMA producer + TMEM allocator, 17 MMA issuer.
#pragma once
#include "attention_tc.cuh"

namespace b200 {

constexpr int A64_DH = 64;
constexpr int A64_BK = 128;                                 // keys per block
constexpr int A64_THREADS = (2 * ATT_GROUP_WARPS + 2) * 32;
constexpr int A64_NSLOT = 3;
constexpr int A64_TILE_BYTES = 128 * A64_DH * 2;            // 16 KB: a query tile, a key block or a value block
constexpr int A64_SLOT_BYTES = 2 * A64_TILE_BYTES;          // K_b then V_b
constexpr int A64_ON = A64_DH + 16;                         // N of the output MMAs
constexpr int A64_Q_OFF = 0;                                // 4 x 16 KB: [group][2] query tiles
constexpr int A64_KV_OFF = A64_Q_OFF + 4 * A64_TILE_BYTES;  // 3 x 32 KB
constexpr int A64_KX_OFF = A64_KV_OFF + A64_NSLOT * A64_SLOT_BYTES;   // 8 KB: [key][1, 0, ...], 64-byte rows
constexpr int A64_QX_OFF = A64_KX_OFF + 128 * 64;           // 2 x 8 KB: per group [row][-m_r, 0, ...]
constexpr int A64_VX_OFF = A64_QX_OFF + 2 * 128 * 64;       // 2 KB of bf16 ones: second N atom of every V step
constexpr int A64_OUT_OFF = A64_VX_OFF + 2048;              // 16 x 2 KB: per softmax warp 32 rows x 32 dims
constexpr int A64_TAIL_OFF = A64_OUT_OFF + 2 * ATT_GROUP_WARPS * 2048;
constexpr int A64_TAIL_BYTES = 256 + 2 * 256 * 4;           // barriers + TMEM slot, exact-maximum exchange
constexpr size_t A64_SMEM_BYTES = 1024 + (size_t)A64_TAIL_OFF + A64_TAIL_BYTES;
static_assert(A64_SMEM_BYTES <= 232448, "attention_tc64 shared memory plan exceeds the 227 KB limit");

struct Att64Params {
  int S, H, heads, n_mt, n_units;
  const int32_t* lens;
  const float* kmax2;        // [batch * heads] max_j |k_j|^2 (att_norms_kernel)
  const int32_t* exact_flag; // != 0: every tile takes the exact-maximum script
  int force_exact;           // option att_exact_max
  int reverse;
};

// max_j |k_j|^2 and max_r |q_r|^2 of every (sequence, head); raises *flag when the Cauchy-Schwarz reference could
// leave a row's leading probabilities outside the normal fp32 / bf16 range (see ATT_BOUND_GAP).
template <int DH>
__global__ void __launch_bounds__(256) att_norms_kernel(const __nv_bfloat16* __restrict__ qkv, int S, int H, int heads,
                                                        float* __restrict__ kmax2, int32_t* __restrict__ flag) {
  // DH / 8 lanes per row, 16 bytes each: a row's 2 * DH bytes are one contiguous, fully used segment
  constexpr int LPR = DH / 8, RPW = 32 / LPR;
  const int h = blockIdx.x, b = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int sub = lane / LPR, c = lane % LPR;
  float q2m = 0.f, k2m = 0.f;
  for (int j0 = warp * RPW; j0 < S; j0 += nwarp * RPW) {
    const int j = j0 + sub;
    float q2 = 0.f, k2 = 0.f;
    if (j < S) {
      const __nv_bfloat16* row = qkv + (size_t)(b * S + j) * 3 * H + h * DH;
      const uint4 wq = __ldg(reinterpret_cast<const uint4*>(row) + c);
      const uint4 wk = __ldg(reinterpret_cast<const uint4*>(row + H) + c);
      const uint32_t a[4] = {wq.x, wq.y, wq.z, wq.w}, kk[4] = {wk.x, wk.y, wk.z, wk.w};
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float q0 = __uint_as_float(a[t] << 16), q1 = __uint_as_float(a[t] & 0xFFFF0000u);
        const float k0 = __uint_as_float(kk[t] << 16), k1 = __uint_as_float(kk[t] & 0xFFFF0000u);
        q2 = fmaf(q0, q0, fmaf(q1, q1, q2));
        k2 = fmaf(k0, k0, fmaf(k1, k1, k2));
      }
    }
#pragma unroll
    for (int o = LPR / 2; o > 0; o >>= 1) {
      q2 += __shfl_xor_sync(0xffffffffu, q2, o);
      k2 += __shfl_xor_sync(0xffffffffu, k2, o);
    }
    q2m = fmaxf(q2m, q2);
    k2m = fmaxf(k2m, k2);
  }
  __shared__ float sq[8], sk[8];
  q2m = warp_max(q2m);
  k2m = warp_max(k2m);
  if (lane == 0) { sq[warp] = q2m; sk[warp] = k2m; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < nwarp; ++w) { q2m = fmaxf(q2m, sq[w]); k2m = fmaxf(k2m, sk[w]); }
    kmax2[b * heads + h] = k2m;
    const float gap_sq = 0.25f * ATT_BOUND_GAP * ATT_BOUND_GAP / (ATT_NORM_SLACK * ATT_NORM_SLACK);
    if (!(q2m * k2m <= gap_sq)) atomicOr(flag, 1);
  }
}

// The same decision from maxima the QKV projection's epilogue has already folded (EPI_HEADNORM): one thread per
// (sequence, head).
__global__ void att_flag_kernel(const float* __restrict__ qmax2, const float* __restrict__ kmax2, int n, int32_t* __restrict__ flag) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float gap_sq = 0.25f * ATT_BOUND_GAP * ATT_BOUND_GAP / (ATT_NORM_SLACK * ATT_NORM_SLACK);
  if (!(qmax2[i] * kmax2[i] <= gap_sq)) atomicOr(flag, 1);
}

struct A64Item {
  int b, h, mt, len, nb, ns;   // sequence, head, query tile, valid keys, key blocks with a valid key, block steps
};

B200_DEVINL A64Item a64_item(const Att64Params& p, long long u, bool exact) {
  A64Item it;
  const int ur = p.reverse ? p.n_units - 1 - (int)u : (int)u;
  const int item = ur / p.n_mt;
  it.mt = ur - item * p.n_mt;
  it.b = item / p.heads;
  it.h = item - it.b * p.heads;
  it.len = max(1, min(p.lens[it.b], p.S));
  it.nb = (it.len + A64_BK - 1) / A64_BK;
  it.ns = exact ? 2 * it.nb : it.nb;
  return it;
}

template <int NPOLY>
__global__ void __launch_bounds__(A64_THREADS, 1)
attention_tc64_kernel(const __grid_constant__ CUtensorMap map_qkv, const __grid_constant__ CUtensorMap map_out,
                      const Att64Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* kx = smem + A64_KX_OFF;
  uint8_t* qx = smem + A64_QX_OFF;
  uint8_t* vx = smem + A64_VX_OFF;
  uint8_t* tail = smem + A64_TAIL_OFF;
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail);
  uint64_t* q_full = bars;            // [4] TMA -> issuer, groups: query tile [group][slot] landed
  uint64_t* q_empty = bars + 4;       // [4] issuer -> TMA: every MMA that reads the tile has finished
  uint64_t* kv_full = bars + 8;       // [3] TMA -> issuer
  uint64_t* kv_empty = bars + 11;     // [3] issuer -> TMA
  uint64_t* s_full = bars + 14;       // [2] issuer -> group: scores of the next block step are in TMEM
  uint64_t* p_ready = bars + 16;      // [2] group -> issuer: probabilities written (or, first pass of exact mode: scores read)
  uint64_t* pv_done = bars + 18;      // [2] issuer -> issuer: the P V product has finished reading the score columns
  uint64_t* o_full = bars + 20;       // [2] issuer -> group: output tile complete
  uint64_t* t_free = bars + 22;       // [2] group -> issuer: output drained
  uint64_t* qx_ready = bars + 24;     // [2] group -> issuer: -m_r of the group's next tile is written
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 26);
  float* xch = reinterpret_cast<float*>(tail + 256);

  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  const int S = p.S;
  constexpr int W_TMA = 2 * ATT_GROUP_WARPS, W_MMA = W_TMA + 1;

  // constant operand slices: kx zeros except [key][0] = 1, qx zeros (rows written per tile), vx all ones
  for (int i = threadIdx.x; i < (128 * 64 * 3) / 16; i += A64_THREADS) reinterpret_cast<uint4*>(kx)[i] = make_uint4(0, 0, 0, 0);
  for (int i = threadIdx.x; i < 2048 / 4; i += A64_THREADS) reinterpret_cast<uint32_t*>(vx)[i] = 0x3F803F80u;
  __syncthreads();
  for (int j = threadIdx.x; j < 128; j += A64_THREADS)      // 64-byte swizzle: 16-byte chunk ^ (row / 2) % 4
    *reinterpret_cast<uint16_t*>(kx + j * 64 + (((j >> 1) & 3) << 4)) = 0x3F80u;
  fence_proxy_async();

  if (warp == W_TMA && lane == 0) {
    tma_prefetch_desc(&map_qkv);
    tma_prefetch_desc(&map_out);
  }
  if (warp == W_MMA && lane == 0) {
    for (int i = 0; i < 4; ++i) { mbar_init(&q_full[i], 1); mbar_init(&q_empty[i], 1); }
    for (int i = 0; i < A64_NSLOT; ++i) { mbar_init(&kv_full[i], 1); mbar_init(&kv_empty[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_ready[i], ATT_GROUP_WARPS);
      mbar_init(&pv_done[i], 1);
      mbar_init(&o_full[i], 1);
      mbar_init(&t_free[i], ATT_GROUP_WARPS);
      mbar_init(&qx_ready[i], ATT_GROUP_WARPS);
    }
    fence_mbar_init();
  }
  if (warp == W_TMA) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                                  // the QKV projection (and att_norms_kernel) have finished
  pdl_launch_dependents();
  const bool exact = p.force_exact || (__ldg(p.exact_flag) != 0);
  const long long stride = gridDim.x;

  if (warp == W_TMA) {
    // ===== TMA producer: query tiles and the K_b / V_b ring, in the issuer's order =====
    uint32_t n_kv = 0;
    for (uint32_t pp = 0;; ++pp) {
      const long long u0 = (long long)blockIdx.x + (long long)(2 * pp) * stride;
      if (u0 >= p.n_units) break;
      A64Item item[2];
      bool have[2] = {true, u0 + stride < p.n_units};
      item[0] = a64_item(p, u0, exact);
      item[1] = have[1] ? a64_item(p, u0 + stride, exact) : item[0];
      const int j = pp & 1;
      for (int g = 0; g < 2; ++g) {
        if (!have[g]) continue;
        mbar_wait(&q_empty[g * 2 + j], ((pp >> 1) & 1) ^ 1);
        if (elect_one()) {
          mbar_arrive_expect_tx(&q_full[g * 2 + j], A64_TILE_BYTES);
          tma_load_2d(smem + A64_Q_OFF + (g * 2 + j) * A64_TILE_BYTES, &map_qkv, &q_full[g * 2 + j], item[g].h * A64_DH,
                      item[g].b * S + item[g].mt * 128);
        }
        __syncwarp();
      }
      const int smax = max(item[0].ns, have[1] ? item[1].ns : 0);
      for (int s = 0; s < smax; ++s) {
        for (int g = 0; g < 2; ++g) {
          if (!have[g] || s >= item[g].ns) continue;
          const bool phase1 = exact && s < item[g].nb;
          const int blk = s % item[g].nb;
          const int slot = n_kv % A64_NSLOT;
          mbar_wait(&kv_empty[slot], ((n_kv / A64_NSLOT) & 1) ^ 1);
          if (elect_one()) {
            uint8_t* base = smem + A64_KV_OFF + slot * A64_SLOT_BYTES;
            mbar_arrive_expect_tx(&kv_full[slot], phase1 ? A64_TILE_BYTES : A64_SLOT_BYTES);
            tma_load_2d(base, &map_qkv, &kv_full[slot], p.H + item[g].h * A64_DH, item[g].b * S + blk * A64_BK);
            if (!phase1)
              tma_load_2d(base + A64_TILE_BYTES, &map_qkv, &kv_full[slot], 2 * p.H + item[g].h * A64_DH, item[g].b * S + blk * A64_BK);
          }
          __syncwarp();
          ++n_kv;
        }
      }
    }
  } else if (warp == W_MMA) {
    // ===== MMA issuer: the two groups in lock step, block by block =====
    const uint32_t idesc_s = umma_idesc_bf16_f32(128, A64_BK);
    const uint32_t idesc_o = umma_idesc_bf16_f32_bmn(128, A64_ON);
    const uint64_t dkx = umma_smem_desc(smem_u32(kx), 4, 512, 0);
    uint32_t n_kv = 0, n_p[2] = {0, 0}, n_pv[2] = {0, 0}, k_g[2] = {0, 0};
    int slot_g[2] = {0, 0};
    // P V of block step s of group g: 16 keys per step, steps of padding only are skipped.  V as an MN-major B operand:
    // a step = 16 key rows of 128 bytes = 2048 bytes; its second N atom (leading-dimension offset) is the block of ones.
    auto issue_pv = [&](int g, const A64Item& it, int s) {
      const int blk = s % it.nb;
      const int vk = min(A64_BK, it.len - blk * A64_BK);
      const int ksteps = (vk + 15) >> 4;
      const uint32_t tb = tmem_base + g * 256;
      const uint32_t vbase = smem_u32(smem + A64_KV_OFF + slot_g[g] * A64_SLOT_BYTES + A64_TILE_BYTES);
      uint64_t dv = umma_smem_desc(vbase, 2, 1024, smem_u32(vx) - vbase);
      if (elect_one()) {
        for (int k = 0; k < ksteps; ++k) {
          const uint32_t tp = tb + (k < 4 ? 8 * k : 64 + 8 * (k - 4));
          umma_bf16_ts(tb + A64_BK, tp, dv, idesc_o, (uint32_t)((k | blk) != 0));
          dv += 128ull - (128ull << 16);           // start + 2048 bytes, leading offset - 2048 bytes
        }
      }
      __syncwarp();
    };
    for (uint32_t pp = 0;; ++pp) {
      const long long u0 = (long long)blockIdx.x + (long long)(2 * pp) * stride;
      if (u0 >= p.n_units) break;
      A64Item item[2];
      bool have[2] = {true, u0 + stride < p.n_units};
      item[0] = a64_item(p, u0, exact);
      item[1] = have[1] ? a64_item(p, u0 + stride, exact) : item[0];
      const int j = pp & 1;
      const int smax = max(item[0].ns, have[1] ? item[1].ns : 0);
      for (int s = 0; s < smax; ++s) {
        for (int g = 0; g < 2; ++g) {
          if (!have[g] || s >= item[g].ns) continue;
          const bool phase1 = exact && s < item[g].nb;
          if (s == 0) {
            mbar_wait(&t_free[g], (k_g[g] & 1) ^ 1);        // the group's previous output has been drained
            mbar_wait(&q_full[g * 2 + j], (pp >> 1) & 1);
            mbar_wait(&qx_ready[g], k_g[g] & 1);
          } else {
            mbar_wait(&p_ready[g], n_p[g] & 1);
            ++n_p[g];
            tc_fence_after();
            if (!(exact && s - 1 < item[g].nb)) {
              issue_pv(g, item[g], s - 1);
              if (elect_one()) {
                umma_commit(&pv_done[g]);
                umma_commit(&kv_empty[slot_g[g]]);
              }
              __syncwarp();
              mbar_wait(&pv_done[g], n_pv[g] & 1);           // the next scores overwrite the probabilities just consumed
              ++n_pv[g];
            }
          }
          const int slot = n_kv % A64_NSLOT;
          mbar_wait(&kv_full[slot], (n_kv / A64_NSLOT) & 1);
          ++n_kv;
          tc_fence_after();
          const uint32_t qbase = smem_u32(smem + A64_Q_OFF + (g * 2 + j) * A64_TILE_BYTES);
          const uint32_t kbase = smem_u32(smem + A64_KV_OFF + slot * A64_SLOT_BYTES);
          if (elect_one()) {
            const uint64_t dq = umma_smem_desc_sw128(qbase), dk = umma_smem_desc_sw128(kbase);
#pragma unroll
            for (int k = 0; k < A64_DH / 16; ++k)
              umma_bf16(tmem_base + g * 256, dq + (uint64_t)(k * 2), dk + (uint64_t)(k * 2), idesc_s, (uint32_t)(k != 0));
            umma_bf16(tmem_base + g * 256, umma_smem_desc(smem_u32(qx + g * 128 * 64), 4, 512, 0), dkx, idesc_s, 1u);
            umma_commit(&s_full[g]);
            if (phase1) umma_commit(&kv_empty[slot]);
          }
          __syncwarp();
          if (!phase1) slot_g[g] = slot;
        }
      }
      for (int g = 0; g < 2; ++g) {
        if (!have[g]) continue;
        mbar_wait(&p_ready[g], n_p[g] & 1);
        ++n_p[g];
        tc_fence_after();
        issue_pv(g, item[g], item[g].ns - 1);
        if (elect_one()) {
          umma_commit(&o_full[g]);
          umma_commit(&kv_empty[slot_g[g]]);
          umma_commit(&q_empty[g * 2 + j]);
        }
        __syncwarp();
        ++k_g[g];
      }
    }
  } else {
    // ===== softmax + output: two threads per query row, one per key half of every block =====
    const int grp = warp / ATT_GROUP_WARPS, w8 = warp % ATT_GROUP_WARPS;
    const int wq = w8 & 3, half = w8 >> 2;
    const int r = wq * 32 + lane;                              // row within the query tile = TMEM lane
    const uint32_t lane_base = tmem_base + grp * 256 + ((uint32_t)(wq * 32) << 16);
    const uint32_t tS = lane_base + half * 64;                 // this half's 64 score columns of a block
    const uint32_t tO = lane_base + A64_BK + half * 32;        // this thread's 32 of the 64 output dims
    const uint32_t tL = lane_base + A64_BK + A64_DH;           // the row sum
    const int pair_bar = 1 + grp * 4 + wq;                     // named barrier of the two warps sharing rows
    const uint32_t s_xm = smem_u32(xch) + grp * 256 * 4;       // [2 halves][128] exact row maxima
    uint8_t* out_w = smem + A64_OUT_OFF + warp * 2048;         // this warp's 32 rows x 64 bytes
    const uint32_t s_out = smem_u32(out_w) + lane * 64;
    const uint32_t s_qx_row = smem_u32(qx) + grp * 128 * 64 + r * 64 + (((r >> 1) & 3) << 4);   // element 0 of row r

    // Softmax reference of a tile: |q_r| from the query tile in shared memory (128-byte swizzle), max |k| from
    // att_norms_kernel; -m_r (or 0 in exact mode) into the A slice of the extra K-step.
    auto reference = [&](uint32_t pp, const A64Item& it) {
      const int j = pp & 1;
      mbar_wait(&q_full[grp * 2 + j], (pp >> 1) & 1);
      if (half == 0) {
        float ms = 0.f;
        if (!exact) {
          const uint32_t qr = smem_u32(smem + A64_Q_OFF + (grp * 2 + j) * A64_TILE_BYTES) + r * 128;
          const int sw = r & 7;
          __nv_bfloat162 a0 = __float2bfloat162_rn(0.f), a1 = a0;
#pragma unroll
          for (int g = 0; g < 8; ++g) att_sumsq8(lds128(qr + ((g ^ sw) << 4)), a0, a1);
          const float q2 = att_sumsq_total(a0, a1);
          ms = att_sqrt_fma(q2 * __ldg(p.kmax2 + it.b * p.heads + it.h)) * ATT_NORM_SLACK;
        }
        sts16(s_qx_row, __bfloat16_as_ushort(__float2bfloat16_rn(-ms)));
        fence_proxy_async();
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&qx_ready[grp]);
    };

    uint32_t n_s = 0, k = 0;
    long long u = (long long)blockIdx.x + (long long)grp * stride;
    bool have = u < p.n_units;
    A64Item it = a64_item(p, have ? u : 0, exact);
    uint32_t pp = 0;
    if (have) reference(pp, it);
    while (have) {
      float mx = -INFINITY;
      for (int s = 0; s < it.ns; ++s) {
        const bool phase1 = exact && s < it.nb;
        const int blk = s % it.nb;
        const int lenh = it.len - blk * A64_BK - half * 64;    // valid keys counted from this half's first
        const bool all_valid = lenh >= 64;
        mbar_wait(&s_full[grp], n_s & 1);
        ++n_s;
        tc_fence_after();
        uint32_t v[2][32];
        if (phase1) {
          const float m = all_valid ? att_pass_max<true>(tS, 2, lenh, v) : att_pass_max<false>(tS, 2, lenh, v);
          mx = fmaxf(mx, m);
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&p_ready[grp]);
          if (s == it.nb - 1) {
            sts32(s_xm + (half * 128 + r) * 4, __float_as_uint(mx));
            named_bar_sync(pair_bar, 64);
            mx = fmaxf(mx, __uint_as_float(lds32(s_xm + ((half ^ 1) * 128 + r) * 4)));   // finite: key 0 is always valid
            named_bar_sync(pair_bar, 64);              // both halves have read before the next tile writes
          }
          continue;
        }
        // probabilities in place: chunk c of this half -> 16 columns from tS + 16 c on
        tmem_ld32(tS, v[0]);
        tmem_ld32(tS + 32, v[1]);
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          tmem_ld_wait_dep(v[c]);
          uint32_t pk[16];
          if (exact) {
            if (all_valid) att_probs<true, true, NPOLY>(v[c], lenh - c * 32, mx, pk);
            else att_probs<false, true, NPOLY>(v[c], lenh - c * 32, mx, pk);
          } else {
            if (all_valid) att_probs<true, false, NPOLY>(v[c], lenh - c * 32, 0.f, pk);
            else att_probs<false, false, NPOLY>(v[c], lenh - c * 32, 0.f, pk);
          }
          tmem_st16(tS + c * 16, pk);
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_ready[grp]);
      }
      // while the last P V product runs: the reference of this group's next tile
      const long long un = u + 2 * stride;
      const bool have_n = un < p.n_units;
      const A64Item it_n = a64_item(p, have_n ? un : 0, exact);
      if (have_n) reference(pp + 1, it_n);
      // output tile: O / l -> bf16 -> this warp's staging rows -> TMA store
      mbar_wait(&o_full[grp], k & 1);
      tc_fence_after();
      uint32_t o[32], lsum;
      tmem_ld32(tO, o);
      tmem_ld1(tL, lsum);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&t_free[grp]);
        tma_store_wait_read();                                 // the previous tile's store has left the staging rows
      }
      __syncwarp();
      {
        const float inv = att_rcp_fma(__uint_as_float(lsum));
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 w;
          __nv_bfloat162 t0 = __floats2bfloat162_rn(__uint_as_float(o[g * 8 + 0]) * inv, __uint_as_float(o[g * 8 + 1]) * inv);
          __nv_bfloat162 t1 = __floats2bfloat162_rn(__uint_as_float(o[g * 8 + 2]) * inv, __uint_as_float(o[g * 8 + 3]) * inv);
          __nv_bfloat162 t2 = __floats2bfloat162_rn(__uint_as_float(o[g * 8 + 4]) * inv, __uint_as_float(o[g * 8 + 5]) * inv);
          __nv_bfloat162 t3 = __floats2bfloat162_rn(__uint_as_float(o[g * 8 + 6]) * inv, __uint_as_float(o[g * 8 + 7]) * inv);
          w.x = *reinterpret_cast<uint32_t*>(&t0); w.y = *reinterpret_cast<uint32_t*>(&t1);
          w.z = *reinterpret_cast<uint32_t*>(&t2); w.w = *reinterpret_cast<uint32_t*>(&t3);
          sts128(s_out + g * 16, w);
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) {
        tma_store_2d(&map_out, out_w, it.h * A64_DH + half * 32, it.b * S + it.mt * 128 + wq * 32);
        tma_store_commit();
      }
      it = it_n;
      u = un;
      have = have_n;
      ++pp;
      ++k;
    }
    if (lane == 0) tma_store_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == W_TMA) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace b200
