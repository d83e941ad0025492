// attention_tc.cuh -- K3 on tcgen05: softmax(QK^T/sqrt(dh) + key-padding mask) V for 32-wide heads
// and sequences up to 256 tokens (the MiniLM-L6 shape).
//
// Work unit = (sequence, head, 128-row query tile).  Per unit:
//   scores   S = Q K^T     one 128 x S tile (M128 x N<=256 x K16, two K steps), fp32 in TMEM
//   softmax  TWO threads per query row (TMEM lane), each owning half of the key chunks: partial row
//            maxima and sums meet through shared memory; P = 2^((s - max) * scale * log2 e) is written
//            back OVER the scores as bf16 (two per 32-bit column) with tcgen05.st
//   output   O = P V       A operand straight from TMEM (no shared-memory round trip for P), V read
//            as an MN-major B operand exactly as TMA lands it ([key][dim] rows of 64 bytes)
//
// One CTA per SM owns all 512 TMEM columns as two score buffers and alternates its units between two
// softmax warp groups.  Warp roles: 0-7 group 0, 8-15 group 1 (within a group warp w: TMEM lane group
// w & 3, key half w >> 2), 16 TMA producer (Q, K, V of the units ahead, four stages) and TMEM
// allocator, 17 MMA issuer.
//
// The tensor work is ~1/8 of the kernel; what bounds it is the exponential (one SFU op per score, 16
// lanes per clock per SM = 2048 clocks per unit).  Two things keep the SFU queue full: (1) two warps
// per scheduler inside the exponential pass (one cannot cover its own TMEM loads and stores), and (2)
// the groups take TURNS in that pass -- a token (mbarrier) handed over at its end -- so that one
// group's maximum pass, output drain and MMA round trips always run under the other's exponentials.
// Left to themselves two independent groups (or two CTAs per SM) fall into lock-step, sharing the SFU
// during the pass and idling it together afterwards (measured: 52 % SFU utilisation).
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int ATT_DH = 32;
constexpr int ATT_MAX_S = 256;
constexpr int ATT_GROUP_WARPS = 8;                       // softmax warps per group
constexpr int ATT_THREADS = (2 * ATT_GROUP_WARPS + 2) * 32;
constexpr int ATT_STAGES = 4;
#ifndef ATT_NG
#define ATT_NG 2
#endif
constexpr int ATT_Q_BYTES = 128 * ATT_DH * 2;            // 8 KB: one 128-row query tile
constexpr int ATT_KV_BYTES = ATT_MAX_S * ATT_DH * 2;     // 16 KB: K or V of one (sequence, head)
constexpr int ATT_STAGE_BYTES = ATT_Q_BYTES + 2 * ATT_KV_BYTES;
constexpr int ATT_TMEM_COLS = 512;                       // two score buffers of 256 columns
constexpr int ATT_OUT_BYTES = 128 * ATT_DH * 2;          // 8 KB per group: output tile staged for its TMA store
constexpr int ATT_SMEM_TAIL = 256 + 2 * 4 * 128 * 4;     // barriers + unit records, per group row max / sum exchange
constexpr size_t ATT_SMEM_BYTES = 1024 + ATT_STAGES * (size_t)ATT_STAGE_BYTES + 2 * ATT_OUT_BYTES + ATT_SMEM_TAIL;

struct AttParams {
  int S;            // tokens per sequence (multiple of 32, <= 256)
  int H;            // hidden size
  int heads;
  int n_mt;         // 128-row query tiles per sequence
  int n_units;      // batch * heads * n_mt
  const int32_t* lens;
  __nv_bfloat16* ctx;
};

#ifdef B200_ATT_TIMING
// Development aid (not compiled into the shipped library): per-phase clock totals of the softmax
// warps, accumulated by one lane per CTA.  [0] wait for scores, [1] pass 1, [2] pass 2, [3] wait for
// output, [4] output, [5] units.
__device__ unsigned long long g_att_timing[16];
#define ATT_STAMP(var) const long long var = clock64()
#define ATT_ACC(i, a, b) do { if (warp == 0 && lane == 0) atomicAdd(&g_att_timing[i], (unsigned long long)((b) - (a))); } while (0)
#else
#define ATT_STAMP(var)
#define ATT_ACC(i, a, b)
#endif

B200_DEVINL float att_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Maximum of one 32-key chunk of a row; FULL = every key valid, else only the first `valid` keys
// (-inf when there are none).
template <bool FULL>
B200_DEVINL float att_chunk_max(const uint32_t (&v)[32], int valid) {
  float m0, m1;
  if (FULL) {
    m0 = __uint_as_float(v[0]);
    m1 = __uint_as_float(v[1]);
#pragma unroll
    for (int j = 2; j < 32; j += 2) {
      m0 = fmaxf(m0, __uint_as_float(v[j]));
      m1 = fmaxf(m1, __uint_as_float(v[j + 1]));
    }
  } else {
    m0 = m1 = -INFINITY;
#pragma unroll
    for (int j = 0; j < 32; ++j) m0 = fmaxf(m0, (j < valid) ? __uint_as_float(v[j]) : -INFINITY);
  }
  return fmaxf(m0, m1);
}

// Probabilities of one 32-key chunk of a row: p = 2^(s*scale - ms) for the valid keys, 0 for padding;
// accumulates the row sum and packs bf16 pairs.
template <bool FULL>
B200_DEVINL void att_probs(const uint32_t (&v)[32], int valid, float scale_log2, float ms, float& sum0, float& sum1,
                           uint32_t (&pk)[16]) {
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    float p0 = att_exp2(fmaf(__uint_as_float(v[2 * j]), scale_log2, -ms));
    float p1 = att_exp2(fmaf(__uint_as_float(v[2 * j + 1]), scale_log2, -ms));
    if (!FULL) {
      if (2 * j >= valid) p0 = 0.f;
      if (2 * j + 1 >= valid) p1 = 0.f;
    }
    sum0 += p0;
    sum1 += p1;
    __nv_bfloat162 t = __floats2bfloat162_rn(p0, p1);
    pk[j] = *reinterpret_cast<uint32_t*>(&t);
  }
}

// pass 1 of a warp: maximum over its `ncl` chunks starting at TMEM address tS, two loads in flight.
template <bool FULL>
B200_DEVINL float att_pass1(uint32_t tS, int ncl, int lenh, uint32_t (&v)[2][32]) {
  float mx = -INFINITY;
#pragma unroll 1
  for (int c = 0; c < ncl; c += 2) {
    tmem_ld32(tS + c * 32, v[0]);
    if (c + 1 < ncl) tmem_ld32(tS + (c + 1) * 32, v[1]);
    tmem_ld_wait_dep(v[0]);
    mx = fmaxf(mx, att_chunk_max<FULL>(v[0], lenh - c * 32));
    if (c + 1 < ncl) {
      tmem_ld_wait_dep(v[1]);
      mx = fmaxf(mx, att_chunk_max<FULL>(v[1], lenh - (c + 1) * 32));
    }
  }
  return mx;
}

// pass 2 of a warp: probabilities written back in place (16 columns per chunk from tS on); the
// tcgen05.ld of chunk c+1 is in flight while chunk c is processed (two register buffers).  Returns
// the partial row sum.
template <bool FULL>
B200_DEVINL float att_pass2(uint32_t tS, int ncl, int lenh, float scale_log2, float ms, uint32_t (&v)[2][32]) {
  float sum0 = 0.f, sum1 = 0.f;
  if (ncl > 0) tmem_ld32(tS, v[0]);
#pragma unroll 1
  for (int c = 0; c < ncl; c += 2) {
#pragma unroll
    for (int uu = 0; uu < 2; ++uu) {
      const int cc = c + uu;
      if (cc < ncl) {
        tmem_ld_wait_dep(v[uu]);
        if (cc + 1 < ncl) tmem_ld32(tS + (cc + 1) * 32, v[uu ^ 1]);
        uint32_t pk[16];
        att_probs<FULL>(v[uu], lenh - cc * 32, scale_log2, ms, sum0, sum1, pk);
        tmem_st16(tS + cc * 16, pk);
      }
    }
  }
  return sum0 + sum1;
}

// TMEM column plan of a unit (S keys = nch chunks of 32; the first c0 = ceil(nch/2) chunks belong to
// key half 0, the rest to half 1):
//   scores           chunk c in columns [32c, 32c + 32)
//   probabilities    in place, at the start of the OWNING half's score columns, 16 columns per chunk:
//                    half 0 chunk c -> [16c, ..), half 1 chunk c -> [32 c0 + 16 (c - c0), ..)
//                    (never ahead of the scores that warp still has to read)
//   output (fp32)    32 columns that hold neither: right after the scores when S <= 224, else the
//                    consumed upper part of half 0's scores, [64, 96)
B200_DEVINL int att_prob_col(int c, int c0) { return c < c0 ? 16 * c : 32 * c0 + 16 * (c - c0); }
B200_DEVINL int att_out_col(int S) { return S <= ATT_MAX_S - 32 ? S : 64; }


__global__ void __launch_bounds__(ATT_THREADS, 1)
attention_tc_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_kv,
                    const __grid_constant__ CUtensorMap map_out, const __grid_constant__ CUtensorMap map_out_tail,
                    const AttParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* out_stage = smem + ATT_STAGES * ATT_STAGE_BYTES;   // [2] per group, 64-byte rows, 64-byte swizzle
  uint8_t* tail = out_stage + 2 * ATT_OUT_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail);
  uint64_t* full = bars;             // [4] TMA -> MMA, softmax: Q, K, V and the unit record of a unit landed
  uint64_t* kv_empty = bars + 4;     // [4] MMA -> TMA: the unit's MMAs have read the stage
  uint64_t* s_full = bars + 8;       // [2] MMA -> group: scores ready
  uint64_t* p_ready = bars + 10;     // [2] group -> MMA: probabilities written to TMEM
  uint64_t* o_full = bars + 12;      // [2] MMA -> group: output tile ready
  uint64_t* t_free = bars + 14;      // [2] group -> MMA: the group's TMEM buffer is reusable
  uint64_t* xu_turn = bars + 16;     // [2] other group -> group: your turn in the exponential pass
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 18);
  int4* meta = reinterpret_cast<int4*>(tail + 160);   // [4] per stage: {token row base, head, query tile, length}
  float* xch = reinterpret_cast<float*>(tail + 256);  // per group: [2][128] partial row maxima, [2][128] partial sums

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = p.S;
  const int q_rows = min(S, 128);
  const int nch = S / 32, c0 = (nch + 1) / 2;
  const int o_col = att_out_col(S);
  constexpr int W_TMA = 2 * ATT_GROUP_WARPS, W_MMA = W_TMA + 1;

  if (warp == W_TMA && lane == 0) {
    tma_prefetch_desc(&map_q);
    tma_prefetch_desc(&map_kv);
    tma_prefetch_desc(&map_out);
    tma_prefetch_desc(&map_out_tail);
  }
  if (warp == W_MMA && lane == 0) {
    for (int i = 0; i < ATT_STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_ready[i], ATT_GROUP_WARPS);    // one arrival per softmax warp of the group
      mbar_init(&o_full[i], 1);
      mbar_init(&t_free[i], ATT_GROUP_WARPS);
      mbar_init(&xu_turn[i], ATT_GROUP_WARPS);
    }
    fence_mbar_init();
  }
  if (warp == W_TMA) tmem_alloc(tmem_slot, ATT_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t kv_bytes = (uint32_t)S * ATT_DH * 2, q_bytes = (uint32_t)q_rows * ATT_DH * 2;

  if (warp == W_TMA) {
    // ===== TMA producer =====
    if (lane == 0) {
      uint32_t it = 0;
      for (int u = blockIdx.x; u < p.n_units; u += gridDim.x, ++it) {
        const int st = it % ATT_STAGES, item = u / p.n_mt, mt = u - item * p.n_mt;
        const int b = item / p.heads, h = item - b * p.heads;
        const int len = p.lens[b];
        mbar_wait(&kv_empty[st], ((it / ATT_STAGES) & 1) ^ 1);
        uint8_t* base = smem + st * ATT_STAGE_BYTES;
        meta[st] = make_int4(b * S, h, mt, max(1, min(len, S)));   // published by the arrive below
        mbar_arrive_expect_tx(&full[st], q_bytes + 2 * kv_bytes);
        tma_load_2d(base, &map_q, &full[st], h * ATT_DH, b * S + mt * 128);                                  // Q tile
        tma_load_2d(base + ATT_Q_BYTES, &map_kv, &full[st], p.H + h * ATT_DH, b * S);                        // K
        tma_load_2d(base + ATT_Q_BYTES + ATT_KV_BYTES, &map_kv, &full[st], 2 * p.H + h * ATT_DH, b * S);     // V
      }
    }
  } else if (warp == W_MMA) {
    // ===== MMA issuer: scores of unit j, then P V of unit j-1 (the other group's) =====
    if (lane == 0) {
      const uint32_t idesc_s = umma_idesc_bf16_f32(128, (uint32_t)S);        // scores: N = S keys
      const uint32_t idesc_o = umma_idesc_bf16_f32_bmn(128, ATT_DH);         // output: N = 32 dims, V MN-major
      int len_prev = 0;
      auto issue_pv = [&](uint32_t j, int len) {
        const int g = j % ATT_NG, st = j % ATT_STAGES;
        const uint32_t tb = tmem_base + g * 256;
        const uint32_t sbase = smem_u32(smem + st * ATT_STAGE_BYTES);
        mbar_wait(&p_ready[g], (j / ATT_NG) & 1);
        tc_fence_after();
#ifdef B200_ATT_TIMING
        const long long m0 = clock64();
#endif
        // 16 keys per step; steps made of padding only (probabilities all zero) are skipped
        const int ksteps = (len + 15) >> 4;
        // V as an MN-major B operand: 16 keys per step = 16 rows of 64 bytes = 1024 bytes (+64 in the
        // descriptor's 16-byte units); probabilities: 8 columns per step, with a jump where key half 1 starts
        uint64_t dv = umma_smem_desc(sbase + ATT_Q_BYTES + ATT_KV_BYTES, 4, 512, 16384);
        uint32_t tp = tb;
        const int k_half = 2 * c0;
        for (int k = 0; k < ksteps; ++k) {
          if (k == k_half) tp = tb + 32 * c0;
          umma_bf16_ts(tb + o_col, tp, dv, idesc_o, (uint32_t)(k != 0));
          dv += 64;
          tp += 8;
        }
        umma_commit(&o_full[g]);
        umma_commit(&kv_empty[st]);   // the stage is free once every MMA issued so far has finished
#ifdef B200_ATT_TIMING
        atomicAdd(&g_att_timing[13], (unsigned long long)(clock64() - m0));
#endif
      };
      uint32_t it = 0;
      for (int u = blockIdx.x; u < p.n_units; u += gridDim.x, ++it) {
        const int g = it % ATT_NG, st = it % ATT_STAGES;
        const uint32_t sbase = smem_u32(smem + st * ATT_STAGE_BYTES);
        mbar_wait(&full[st], (it / ATT_STAGES) & 1);
        const int len = meta[st].w;
        mbar_wait(&t_free[g], ((it / ATT_NG) & 1) ^ 1);  // the group's previous output has been drained
        tc_fence_after();
        // Q (A) and K (B) are K-major: rows of 64 bytes, 8-row swizzle atoms of 512 bytes
#pragma unroll
        for (int k = 0; k < ATT_DH / 16; ++k) {
          const uint64_t dq = umma_smem_desc(sbase + k * 32, 4, 512, 0);
          const uint64_t dk = umma_smem_desc(sbase + ATT_Q_BYTES + k * 32, 4, 512, 0);
          umma_bf16(tmem_base + g * 256, dq, dk, idesc_s, (uint32_t)(k != 0));
        }
        umma_commit(&s_full[g]);
        if (ATT_NG == 1) issue_pv(it, len);
        else if (it > 0) issue_pv(it - 1, len_prev);
        len_prev = len;
      }
      if (ATT_NG == 2 && it > 0) issue_pv(it - 1, len_prev);
    }
  } else {
    // ===== softmax + output: two threads per query row, one per key half =====
    const int grp = warp / ATT_GROUP_WARPS, w8 = warp % ATT_GROUP_WARPS;
    const int wq = w8 & 3, half = w8 >> 2;
    const int cbeg = half ? c0 : 0;                            // this warp's chunks: [cbeg, cbeg + ncl)
    const int ncl = half ? nch - c0 : c0;
    const int r = wq * 32 + lane;                              // row within the query tile = TMEM lane
    const uint32_t lane_base = tmem_base + grp * 256 + ((uint32_t)(wq * 32) << 16);
    const uint32_t tS = lane_base + cbeg * 32;                 // first score column of this key half
    const uint32_t tO = lane_base + o_col + half * 16;         // this thread's 16 of the 32 output dims
    const float scale_log2 = rsqrtf((float)ATT_DH) * 1.4426950408889634f;
    const int pair_bar = 1 + grp * 4 + wq;                     // named barrier of the two warps sharing rows
    const int group_bar = 9 + grp;                             // named barrier of the whole group
    float* xmax = xch + grp * 512;
    float* xsum = xmax + 256;
    uint32_t k = 0;                                            // units this group has done
    for (uint32_t it = grp; grp < ATT_NG && (long long)blockIdx.x + (long long)it * gridDim.x < p.n_units; it += ATT_NG, ++k) {
      const int st = it % ATT_STAGES;
      ATT_STAMP(t0);
      mbar_wait(&full[st], (it / ATT_STAGES) & 1);             // the producer's unit record is visible
      const int4 um = meta[st];
      const int lenh = um.w - cbeg * 32;                       // valid keys counted from this half's first
      mbar_wait(&s_full[grp], k & 1);
      tc_fence_after();
      ATT_STAMP(t1);
      uint32_t v[2][32];
      const bool all_valid = lenh >= ncl * 32;                 // no padding among this warp's keys (the usual case)
      // pass 1: maximum over this half's valid keys
      float mx = all_valid ? att_pass1<true>(tS, ncl, lenh, v) : att_pass1<false>(tS, ncl, lenh, v);
      xmax[half * 128 + r] = mx;
      named_bar_sync(pair_bar, 64);
      mx = fmaxf(mx, xmax[(half ^ 1) * 128 + r]);              // finite: key 0 is always valid
      const float ms = mx * scale_log2;
      ATT_STAMP(t2);
      // pass 2: probabilities (bf16, packed two per column) over the scores; partial row sum in fp32.
      // Entered only on this group's turn: group 0 goes first, then the groups alternate.
      if (ATT_NG == 2) mbar_wait(&xu_turn[grp], grp == 0 ? ((k & 1) ^ 1) : (k & 1));
      ATT_STAMP(t2b);
      const float psum = all_valid ? att_pass2<true>(tS, ncl, lenh, scale_log2, ms, v)
                                   : att_pass2<false>(tS, ncl, lenh, scale_log2, ms, v);
      __syncwarp();
      if (ATT_NG == 2 && lane == 0) mbar_arrive(&xu_turn[grp ^ 1]);           // hand the exponential pass over
      xsum[half * 128 + r] = psum;
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_ready[grp]);
      // output tile: O / sum -> bf16 -> staged in shared memory -> one TMA store per unit
      ATT_STAMP(t3);
      mbar_wait(&o_full[grp], k & 1);
      tc_fence_after();
      ATT_STAMP(t4);
      uint32_t o[16];
      tmem_ld16(tO, o);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&t_free[grp]);
      if (w8 == 0 && lane == 0) tma_store_wait_read();         // the previous unit's store has left the staging tile
      named_bar_sync(group_bar, ATT_GROUP_WARPS * 32);         // ... and every partial sum of this unit is visible
      {
        float inv;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(xsum[r] + xsum[128 + r]));
        // 64-byte rows, 16-byte chunk index XOR (row / 2) % 4 (the 64-byte TMA swizzle)
        uint8_t* orow = out_stage + grp * ATT_OUT_BYTES + r * 64;
        const int sw = (r >> 1) & 3;
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          uint4 w;
          __nv_bfloat162 t0 = __floats2bfloat162_rn(__uint_as_float(o[g * 8 + 0]) * inv, __uint_as_float(o[g * 8 + 1]) * inv);
          __nv_bfloat162 t1 = __floats2bfloat162_rn(__uint_as_float(o[g * 8 + 2]) * inv, __uint_as_float(o[g * 8 + 3]) * inv);
          __nv_bfloat162 t2 = __floats2bfloat162_rn(__uint_as_float(o[g * 8 + 4]) * inv, __uint_as_float(o[g * 8 + 5]) * inv);
          __nv_bfloat162 t3 = __floats2bfloat162_rn(__uint_as_float(o[g * 8 + 6]) * inv, __uint_as_float(o[g * 8 + 7]) * inv);
          w.x = *reinterpret_cast<uint32_t*>(&t0); w.y = *reinterpret_cast<uint32_t*>(&t1);
          w.z = *reinterpret_cast<uint32_t*>(&t2); w.w = *reinterpret_cast<uint32_t*>(&t3);
          *reinterpret_cast<uint4*>(orow + (((half * 2 + g) ^ sw) << 4)) = w;
        }
      }
      fence_proxy_async();
      named_bar_sync(group_bar, ATT_GROUP_WARPS * 32);
      if (w8 == 0 && lane == 0) {
        // the last query tile of a sequence may be shorter than 128 rows: its own (smaller) store box
        tma_store_2d(um.z == p.n_mt - 1 ? &map_out_tail : &map_out, out_stage + grp * ATT_OUT_BYTES, um.y * ATT_DH,
                     um.x + um.z * 128);
        tma_store_commit();
      }
#ifdef B200_ATT_TIMING
      ATT_STAMP(t5);
      ATT_ACC(0, t0, t1); ATT_ACC(1, t1, t2); ATT_ACC(6, t2, t2b); ATT_ACC(2, t2b, t3); ATT_ACC(3, t3, t4); ATT_ACC(4, t4, t5);
      ATT_ACC(5, 0, 1);
#endif
    }
  }

  if (warp < 2 * ATT_GROUP_WARPS && (warp % ATT_GROUP_WARPS) == 0 && lane == 0) tma_store_wait_all();
  tc_fence_before();
  __syncthreads();
  if (warp == W_TMA) {
    tc_fence_after();
    tmem_dealloc(tmem_base, ATT_TMEM_COLS);
  }
}

}  // namespace b200
