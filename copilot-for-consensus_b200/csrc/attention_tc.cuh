// attention_tc.cuh -- K3 on tcgen05: softmax(QK^T/sqrt(dh) + key-padding mask) V for 32-wide heads
// and sequences up to 256 tokens (the MiniLM-L6 shape).
//
// Work unit = (sequence, head, 128-row query tile).  What bounds a unit is the exponential: one SFU
// op per score, 16 lanes per clock per SM = 2048 clocks; next come the issue slots (a classic softmax
// spends 3.5 instructions per score) and the 71-clock floor of each small tcgen05.mma.  The kernel is
// arranged so that the softmax warps execute little besides the exponential itself:
//
//   scores   S' = [Q | -m] [K | 1]^T : a third K-step whose A slice holds -m_r (the row's softmax
//            reference) and whose B slice holds ones, so the tensor core does the subtraction;
//            log2(e)/sqrt(dh) is folded into the query projection at weight load (encoder.cu), so
//            there is no per-score multiply either.  128 x S fp32 tile in TMEM.
//   m_r      NOT the row maximum (that would be a second pass over the tile, and TMEM reads run at
//            64 B/clk per scheduler): any reference with 2^(s - m) representable gives the same softmax.
//            Cauchy-Schwarz: s_rj <= |q_r| max_j |k_j|, computed from the shared-memory tiles (|q_r| by
//            the row's own threads while the previous unit's P V product runs, max |k| by a helper warp).  Units whose bound is too loose (ATT_BOUND_GAP) fall
//            back to the exact maximum pass.
//   softmax  P = 2^(S') written back OVER the scores as bf16 (two per 32-bit column) with tcgen05.st:
//            per score one MUFU.EX2 and half an F2FP.  TWO threads per query row, one per key half.
//   output   [O | l] = P [V | 1]: A operand straight from TMEM, V read as an MN-major B operand exactly
//            as TMA lands it; a block of ones appended along N yields the row sums l in the same MMAs
//            (no additions in the softmax warps).  O / l -> bf16 -> per-warp staging -> TMA store.
//
// One CTA per SM owns all 512 TMEM columns as two score buffers and alternates its units between two
// softmax warp groups.  Warp roles: 0-7 group 0, 8-15 group 1 (within a group warp w: TMEM lane group
// w & 3, key half w >> 2), 16 TMA producer (Q, K, V of the units ahead, four stages) and TMEM
// allocator, 17 MMA issuer, 18 key norms (max |k|^2 per unit, ahead of the groups).  The groups take TURNS in the exponential pass -- a token (mbarrier) handed
// over at its end -- so that one group's reference computation, output drain and MMA round trips run
// under the other's exponentials.  Left to themselves two groups (or two CTAs per SM) fall into
// lock-step, sharing the SFU during the pass and idling it together afterwards (measured: 52 % SFU
// utilisation).  The group that is not in the pass stays off the SFU altogether (att_rcp_fma,
// att_sqrt_fma).
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int ATT_DH = 32;
constexpr int ATT_MAX_S = 256;
constexpr int ATT_GROUP_WARPS = 8;                       // softmax warps per group
constexpr int ATT_THREADS = (2 * ATT_GROUP_WARPS + 3) * 32;   // + TMA producer, MMA issuer, key-norm warp
constexpr int ATT_STAGES = 4;
constexpr int ATT_Q_BYTES = 128 * ATT_DH * 2;            // 8 KB: one 128-row query tile
constexpr int ATT_KV_BYTES = ATT_MAX_S * ATT_DH * 2;     // 16 KB: K or V of one (sequence, head)
constexpr int ATT_STAGE_BYTES = ATT_Q_BYTES + 2 * ATT_KV_BYTES;
constexpr int ATT_TMEM_COLS = 512;                       // two score buffers of 256 columns
constexpr int ATT_ON = ATT_DH + 16;                      // N of the output MMAs: 32 dims + a block of ones
// shared memory after the stages (offsets from the 1024-aligned end of the last stage)
constexpr int ATT_KX_OFF = 0;                            // 16 KB: B slice of the third K-step, [key][1, 0, ...]
constexpr int ATT_QX_OFF = ATT_KX_OFF + ATT_KV_BYTES;    // 2 x 8 KB: per group A slice of the third K-step, [row][-m_r, 0, ...]
constexpr int ATT_VX_OFF = ATT_QX_OFF + 2 * ATT_Q_BYTES; // 1 KB of bf16 ones: the N block appended to every V step
constexpr int ATT_OUT_OFF = ATT_VX_OFF + 1024;           // 16 x 1 KB: per softmax warp, 32 rows x 16 dims staged for TMA
constexpr int ATT_TAIL_OFF = ATT_OUT_OFF + 2 * ATT_GROUP_WARPS * 1024;
constexpr int ATT_XCH_FLOATS = 256;                      // per group: exact row maxima of the two key halves
constexpr int ATT_TAIL_HDR = 320;                        // barriers, TMEM slot, unit records, per-stage max |k|^2
constexpr int ATT_TAIL_BYTES = ATT_TAIL_HDR + 2 * ATT_XCH_FLOATS * 4;
constexpr size_t ATT_SMEM_BYTES = 1024 + ATT_STAGES * (size_t)ATT_STAGE_BYTES + ATT_TAIL_OFF + ATT_TAIL_BYTES;

struct AttParams {
  int S;            // tokens per sequence (multiple of 32, <= 256)
  int H;            // hidden size
  int heads;
  int n_mt;         // 128-row query tiles per sequence
  int n_units;      // batch * heads * n_mt
  const int32_t* lens;
  int exact_max;    // 1: always take the exact row-maximum pass (B200_ATT_EXACT_MAX=1), for A/B checks
  int reverse;      // 1: walk the units from the last to the first (see GemmParams::reverse)
  int no_turns;     // 1: the groups do not take turns in the exponential pass (B200_ATT_NO_TURNS=1), for A/B timing
};

#ifdef B200_ATT_TIMING
// Development aid (not compiled into the shipped library): per-phase clock totals of group 0's first
// warp.  [0] reference (norms), [1] wait for scores, [2] exact maximum pass, [3] wait for the turn,
// [4] exponential pass, [5] wait for output, [6] output, [7] units.
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

// 2^x on the FMA / ALU pipes: the SFU delivers one MUFU.EX2 per 8 clocks per scheduler and is what bounds this
// kernel, while the FMA pipe issues every clock and sat half idle -- so a fixed share of every chunk's
// exponentials (NPOLY of each 8) is computed here instead and the two pipes run side by side.
//   r = round(x) through the 1.5 * 2^23 trick (its integer lands in the low mantissa bits of t),
//   f = x - r in [-0.5, 0.5],  2^f by a degree-3 polynomial (max relative error 1.0e-4, a twentieth of the bf16
//   rounding step the probability gets next),  2^r by adding r to the exponent field.
// 8 instructions against one MUFU slot of 8 clocks.  x is clamped at -126 (the MUFU path flushes such values to 0,
// this one returns ~1e-38: both vanish against a row's leading probabilities).
B200_DEVINL float att_exp2_fma(float x) {
  x = fmaxf(x, -126.f);
  const float t = x + 12582912.f;
  const float f = x - (t - 12582912.f);
  float p = fmaf(f, 0.055838283f, 0.24263948f);
  p = fmaf(p, f, 0.69313675f);
  p = fmaf(p, f, 0.99992454f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}

// Reciprocal and square root on the FMA / ALU pipes only (bit-trick seed + Newton steps, relative error
// < 1e-6 and < 1e-5): the group that is NOT in its exponential pass must stay off the SFU, whose queue
// the other group keeps full.
B200_DEVINL float att_rcp_fma(float x) {          // x > 0, normal
  float r = __uint_as_float(0x7EF311C7u - __float_as_uint(x));
#pragma unroll
  for (int i = 0; i < 4; ++i) r = r * fmaf(-x, r, 2.0f);
  return r;
}
B200_DEVINL float att_sqrt_fma(float x) {         // x >= 0
  if (x < 1e-30f) return 0.f;
  float y = __uint_as_float(0x5F3759DFu - (__float_as_uint(x) >> 1));      // ~ 1/sqrt(x)
  const float hx = 0.5f * x;
#pragma unroll
  for (int i = 0; i < 3; ++i) y = y * fmaf(-hx * y, y, 1.5f);
  return x * y;
}

// Sum of squares of eight bf16 values packed in a 16-byte word, accumulated as packed bf16 pairs (four
// HFMA2 instead of sixteen conversions and FMAs).  Only ever used for an upper bound, which ATT_NORM_SLACK
// inflates by more than the bf16 accumulation can lose (<= 8 roundings of 2^-9 per accumulator).
B200_DEVINL void att_sumsq8(const uint4 w, __nv_bfloat162& a0, __nv_bfloat162& a1) {
  const uint32_t x[4] = {w.x, w.y, w.z, w.w};
  const __nv_bfloat162 v0 = *reinterpret_cast<const __nv_bfloat162*>(&x[0]), v1 = *reinterpret_cast<const __nv_bfloat162*>(&x[1]);
  const __nv_bfloat162 v2 = *reinterpret_cast<const __nv_bfloat162*>(&x[2]), v3 = *reinterpret_cast<const __nv_bfloat162*>(&x[3]);
  a0 = __hfma2(v0, v0, a0);
  a1 = __hfma2(v1, v1, a1);
  a0 = __hfma2(v2, v2, a0);
  a1 = __hfma2(v3, v3, a1);
}
B200_DEVINL float att_sumsq_total(const __nv_bfloat162 a0, const __nv_bfloat162 a1) {
  const float2 f0 = __bfloat1622float2(a0), f1 = __bfloat1622float2(a1);
  return (f0.x + f0.y) + (f1.x + f1.y);
}
constexpr float ATT_NORM_SLACK = 1.04f;           // on |q| |k|: covers two sums of squares each up to 3.2 % low

// The true row maximum lies at most 2 |q_r| max|k| below the Cauchy-Schwarz reference; while that gap
// (already in log2 units: the scale is folded into q) stays under ATT_BOUND_GAP, every probability of
// the row's leading keys is a normal fp32 / bf16 number and the softmax is unchanged.  Units over the
// limit (decided per unit, uniformly for the group) take the exact maximum pass instead.
constexpr float ATT_BOUND_GAP = 96.f;
#ifndef ATT_HANDOVER
#define ATT_HANDOVER 2   // chunks of its exponential pass a warp still has to do when it hands the turn over (measured: 0 -> 5042, 1 -> 4830, 2 -> 4538, 3 -> 4932, 4 -> 5379 clocks per pair of units)
#endif

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

// Probabilities of one 32-key chunk of a row, packed as bf16 pairs: p = 2^(s') for the valid keys, 0
// for padding.  SUB: the scores carry no reference yet (exact-maximum units), subtract mx first.
template <bool FULL, bool SUB, int NPOLY>
B200_DEVINL void att_probs(const uint32_t (&v)[32], int valid, float mx, uint32_t (&pk)[16]) {
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    float a0 = __uint_as_float(v[2 * j]), a1 = __uint_as_float(v[2 * j + 1]);
    if (SUB) {
      a0 -= mx;
      a1 -= mx;
    }
    // columns c with (c % 8) < NPOLY go to the FMA pipe, spread over the chunk so both pipes stay fed
    float p0 = ((2 * j) % 8 < NPOLY) ? att_exp2_fma(a0) : att_exp2(a0);
    float p1 = ((2 * j + 1) % 8 < NPOLY) ? att_exp2_fma(a1) : att_exp2(a1);
    if (!FULL) {
      if (2 * j >= valid) p0 = 0.f;
      if (2 * j + 1 >= valid) p1 = 0.f;
    }
    __nv_bfloat162 t = __floats2bfloat162_rn(p0, p1);
    pk[j] = *reinterpret_cast<uint32_t*>(&t);
  }
}

// Exact maximum over a warp's `ncl` chunks starting at TMEM address tS, two loads in flight.
template <bool FULL>
B200_DEVINL float att_pass_max(uint32_t tS, int ncl, int lenh, uint32_t (&v)[2][32]) {
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

// Exponential pass of a warp: probabilities written back in place (16 columns per chunk from tS on);
// the tcgen05.ld of chunk c+1 is in flight while chunk c is processed (two register buffers).
// `turn` is handed to the other group when this warp STARTS its last chunk: the overlap of one chunk
// keeps the SFU queue full across the hand-over (warps of a group do not finish together).
template <bool FULL, bool SUB, int NPOLY>
B200_DEVINL void att_pass_exp(uint32_t tS, int ncl, int lenh, float mx, uint32_t (&v)[2][32], uint64_t* turn, int lane) {
  if (ncl > 0) tmem_ld32(tS, v[0]);
  else if (lane == 0) mbar_arrive(turn);
#pragma unroll 1
  for (int c = 0; c < ncl; c += 2) {
#pragma unroll
    for (int uu = 0; uu < 2; ++uu) {
      const int cc = c + uu;
      if (cc < ncl) {
        tmem_ld_wait_dep(v[uu]);
        if (cc + 1 < ncl) tmem_ld32(tS + (cc + 1) * 32, v[uu ^ 1]);
        if (cc == max(ncl - ATT_HANDOVER, 0) && lane == 0) mbar_arrive(turn);
        uint32_t pk[16];
        att_probs<FULL, SUB, NPOLY>(v[uu], lenh - cc * 32, mx, pk);
        tmem_st16(tS + cc * 16, pk);
      }
    }
  }
}

// TMEM column plan of a unit (S keys = nch chunks of 32; the first c0 = ceil(nch/2) chunks belong to
// key half 0, the rest to half 1):
//   scores           chunk c in columns [32c, 32c + 32)
//   probabilities    in place, at the start of the OWNING half's score columns, 16 columns per chunk:
//                    half 0 chunk c -> [16c, ..), half 1 chunk c -> [32 c0 + 16 (c - c0), ..)
//                    (never ahead of the scores that warp still has to read)
//   output (fp32)    ATT_ON = 48 columns (32 dims, then the row sum 16 times) that hold neither: right
//                    after the scores when they fit, else the consumed upper part of half 0's scores
B200_DEVINL int att_prob_col(int c, int c0) { return c < c0 ? 16 * c : 32 * c0 + 16 * (c - c0); }
B200_DEVINL int att_out_col(int S) { return S + ATT_ON <= ATT_MAX_S ? S : 64; }

template <int NPOLY>
__global__ void __launch_bounds__(ATT_THREADS, 1)
attention_tc_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_kv,
                    const __grid_constant__ CUtensorMap map_out, const AttParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* extra = smem + ATT_STAGES * ATT_STAGE_BYTES;
  uint8_t* kx = extra + ATT_KX_OFF;
  uint8_t* qx = extra + ATT_QX_OFF;
  uint8_t* vx = extra + ATT_VX_OFF;
  uint8_t* tail = extra + ATT_TAIL_OFF;
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail);
  uint64_t* full = bars;             // [4] TMA -> MMA, softmax: Q, K, V and the unit record of a unit landed
  uint64_t* kv_empty = bars + 4;     // [4] MMA -> TMA: the unit's MMAs have read the stage
  uint64_t* s_full = bars + 8;       // [2] MMA -> group: scores ready
  uint64_t* p_ready = bars + 10;     // [2] group -> MMA: probabilities written to TMEM
  uint64_t* o_full = bars + 12;      // [2] MMA -> group: output tile ready
  uint64_t* t_free = bars + 14;      // [2] group -> MMA: the group's TMEM buffer is reusable
  uint64_t* xu_turn = bars + 16;     // [2] other group -> group: your turn in the exponential pass
  uint64_t* qx_ready = bars + 18;    // [2] group -> MMA: the reference slice of the group's next unit is written
  uint64_t* kn_ready = bars + 20;    // [4] key-norm warp -> groups: max |k|^2 of the stage's keys is written
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 24);
  int4* meta = reinterpret_cast<int4*>(tail + 208);   // [4] per stage: {token row base, head, query tile, length}
  float* kmax2 = reinterpret_cast<float*>(tail + 272);   // [4] per stage: max_j |k_j|^2
  float* xch = reinterpret_cast<float*>(tail + ATT_TAIL_HDR);

  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  const int S = p.S;
  const int q_rows = min(S, 128);
  const int nch = S / 32, c0 = (nch + 1) / 2;
  const int o_col = att_out_col(S);
  constexpr int W_TMA = 2 * ATT_GROUP_WARPS, W_MMA = W_TMA + 1, W_NORM = W_TMA + 2;

  // constant operand slices: zeros except [key][0] = 1 (kx), all ones (vx); qx rows are written per unit
  for (int i = threadIdx.x; i < (ATT_KV_BYTES + 2 * ATT_Q_BYTES) / 16; i += ATT_THREADS)
    reinterpret_cast<uint4*>(kx)[i] = make_uint4(0, 0, 0, 0);
  for (int i = threadIdx.x; i < 1024 / 4; i += ATT_THREADS) reinterpret_cast<uint32_t*>(vx)[i] = 0x3F803F80u;
  __syncthreads();
  for (int j = threadIdx.x; j < ATT_MAX_S; j += ATT_THREADS)   // 64-byte swizzle: 16-byte chunk ^ (row / 2) % 4
    *reinterpret_cast<uint16_t*>(kx + j * 64 + (((j >> 1) & 3) << 4)) = 0x3F80u;
  fence_proxy_async();

  if (warp == W_TMA && lane == 0) {
    tma_prefetch_desc(&map_q);
    tma_prefetch_desc(&map_kv);
    tma_prefetch_desc(&map_out);
  }
  if (warp == W_MMA && lane == 0) {
    for (int i = 0; i < ATT_STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&kv_empty[i], 1);
      mbar_init(&kn_ready[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_ready[i], ATT_GROUP_WARPS);    // one arrival per softmax warp of the group
      mbar_init(&o_full[i], 1);
      mbar_init(&t_free[i], ATT_GROUP_WARPS);
      mbar_init(&xu_turn[i], ATT_GROUP_WARPS);
      mbar_init(&qx_ready[i], ATT_GROUP_WARPS);
    }
    fence_mbar_init();
  }
  if (warp == W_TMA) tmem_alloc(tmem_slot, ATT_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t kv_bytes = (uint32_t)S * ATT_DH * 2, q_bytes = (uint32_t)q_rows * ATT_DH * 2;
  pdl_wait();                                  // the QKV projection has finished
  pdl_launch_dependents();

  if (warp == W_TMA) {
    // ===== TMA producer: the warp runs the loop converged, one elected lane issues (see uniform_warp_idx) =====
    {
      uint32_t it = 0;
      for (int u = blockIdx.x; u < p.n_units; u += gridDim.x, ++it) {
        const int ur = p.reverse ? p.n_units - 1 - u : u;
        const int st = it % ATT_STAGES, item = ur / p.n_mt, mt = ur - item * p.n_mt;
        const int b = item / p.heads, h = item - b * p.heads;
        const int len = p.lens[b];
        mbar_wait(&kv_empty[st], ((it / ATT_STAGES) & 1) ^ 1);
        uint8_t* base = smem + st * ATT_STAGE_BYTES;
        if (elect_one()) {
          meta[st] = make_int4(b * S, h, mt, max(1, min(len, S)));   // published by the arrive below
          mbar_arrive_expect_tx(&full[st], q_bytes + 2 * kv_bytes);
          tma_load_2d(base, &map_q, &full[st], h * ATT_DH, b * S + mt * 128);                                  // Q tile
          tma_load_2d(base + ATT_Q_BYTES, &map_kv, &full[st], p.H + h * ATT_DH, b * S);                        // K
          tma_load_2d(base + ATT_Q_BYTES + ATT_KV_BYTES, &map_kv, &full[st], 2 * p.H + h * ATT_DH, b * S);     // V
        }
        __syncwarp();
      }
    }
  } else if (warp == W_MMA) {
    // ===== MMA issuer: scores of unit j, then P V of unit j-1 (the other group's).  The warp runs the
    // loop converged and one elected lane issues, which keeps the operands in uniform registers =====
    {
      const uint32_t idesc_s = umma_idesc_bf16_f32(128, (uint32_t)S);        // scores: N = S keys
      const uint32_t idesc_o = umma_idesc_bf16_f32_bmn(128, ATT_ON);         // output: N = 32 dims + 16 ones, V MN-major
      const uint64_t dkx = umma_smem_desc(smem_u32(kx), 4, 512, 0);
      int len_prev = 0;
      auto issue_pv = [&](uint32_t j, int len) {
        const int g = j & 1, st = j % ATT_STAGES;
        const uint32_t tb = tmem_base + g * 256;
        const uint32_t vbase = smem_u32(smem + st * ATT_STAGE_BYTES + ATT_Q_BYTES + ATT_KV_BYTES);
        mbar_wait(&p_ready[g], (j >> 1) & 1);
        tc_fence_after();
        // 16 keys per step; steps made of padding only (probabilities all zero) are skipped.
        // V as an MN-major B operand: a step = 16 key rows of 64 bytes = 1024 bytes; its second N atom
        // (leading-dimension offset) is the block of ones, so the offset shrinks as the start advances.
        const int ksteps = (len + 15) >> 4;
        uint64_t dv = umma_smem_desc(vbase, 4, 512, smem_u32(vx) - vbase);
        uint32_t tp = tb;
        const int k_half = 2 * c0;
        if (elect_one()) {
          for (int k = 0; k < ksteps; ++k) {
            if (k == k_half) tp = tb + 32 * c0;      // probabilities of key half 1 start at its own score columns
            umma_bf16_ts(tb + o_col, tp, dv, idesc_o, (uint32_t)(k != 0));
            dv += 64ull - (64ull << 16);             // start address + 1024 bytes, leading offset - 1024 bytes
            tp += 8;
          }
          umma_commit(&o_full[g]);
          umma_commit(&kv_empty[st]);   // the stage is free once every MMA issued so far has finished
        }
        __syncwarp();
      };
      uint32_t it = 0;
      for (int u = blockIdx.x; u < p.n_units; u += gridDim.x, ++it) {
        const int g = it & 1, st = it % ATT_STAGES;
        const uint32_t sbase = smem_u32(smem + st * ATT_STAGE_BYTES);
        mbar_wait(&full[st], (it / ATT_STAGES) & 1);
        const int len = meta[st].w;
        mbar_wait(&t_free[g], ((it >> 1) & 1) ^ 1);  // the group's previous output has been drained
        mbar_wait(&qx_ready[g], (it >> 1) & 1);      // -m_r of this unit's rows is in the group's A slice
        tc_fence_after();
        // Q (A) and K (B) are K-major: rows of 64 bytes, 8-row swizzle atoms of 512 bytes
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < ATT_DH / 16; ++k) {
            const uint64_t dq = umma_smem_desc(sbase + k * 32, 4, 512, 0);
            const uint64_t dk = umma_smem_desc(sbase + ATT_Q_BYTES + k * 32, 4, 512, 0);
            umma_bf16(tmem_base + g * 256, dq, dk, idesc_s, (uint32_t)(k != 0));
          }
          umma_bf16(tmem_base + g * 256, umma_smem_desc(smem_u32(qx + g * ATT_Q_BYTES), 4, 512, 0), dkx, idesc_s, 1u);
          umma_commit(&s_full[g]);
        }
        __syncwarp();
        if (it > 0) issue_pv(it - 1, len_prev);
        len_prev = len;
      }
      if (it > 0) issue_pv(it - 1, len_prev);
    }
  } else if (warp == W_NORM) {
    // ===== key norms: max_j |k_j|^2 of every unit, one warp running ahead of the softmax groups =====
    const uint32_t s_smem = smem_u32(smem);
    uint32_t it = 0;
    for (int u = blockIdx.x; u < p.n_units; u += gridDim.x, ++it) {
      const int st = it % ATT_STAGES;
      mbar_wait(&full[st], (it / ATT_STAGES) & 1);
      const uint32_t kbase = s_smem + st * ATT_STAGE_BYTES + ATT_Q_BYTES;
      float k2 = 0.f;
      for (int j = lane; j < S; j += 32) {                     // keys lane, lane + 32, ...: 64-byte rows
        const uint32_t kr = kbase + j * 64;
        const int sw = (j >> 1) & 3;                           // visit the 16-byte chunks in a conflict-free order
        __nv_bfloat162 b0 = __float2bfloat162_rn(0.f), b1 = b0;
#pragma unroll
        for (int g = 0; g < 4; ++g) att_sumsq8(lds128(kr + ((g ^ sw) << 4)), b0, b1);
        k2 = fmaxf(k2, att_sumsq_total(b0, b1));
      }
      // non-negative floats order like their bit patterns: one REDUX for the warp maximum
      const uint32_t k2w = __reduce_max_sync(0xffffffffu, __float_as_uint(k2));
      if (lane == 0) {
        kmax2[st] = __uint_as_float(k2w);                      // the stage is not reused before the unit's P V is done
        mbar_arrive(&kn_ready[st]);
      }
      __syncwarp();
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
    const uint32_t tL = lane_base + o_col + ATT_DH;            // the row sum
    const int pair_bar = 1 + grp * 4 + wq;                     // named barrier of the two warps sharing rows
    // shared-space addresses (see lds128): stages, unit records, exchange buffers, staging rows
    const uint32_t s_smem = smem_u32(smem), s_meta = smem_u32(meta);
    const uint32_t s_xm = smem_u32(xch) + grp * ATT_XCH_FLOATS * 4;   // [2 halves][128] exact row maxima
    const uint32_t s_kmax2 = smem_u32(kmax2);
    uint8_t* out_w = extra + ATT_OUT_OFF + warp * 1024;        // this warp's 32 rows x 32 bytes
    const uint32_t s_out = smem_u32(out_w) + lane * 32;
    const uint32_t s_qx_row = smem_u32(qx) + grp * ATT_Q_BYTES + r * 64 + (((r >> 1) & 3) << 4);   // element 0 of row r
    // 2 |q| |k| <= gap, with the norms possibly ATT_NORM_SLACK low
    const float gap_sq = 0.25f * ATT_BOUND_GAP * ATT_BOUND_GAP / (ATT_NORM_SLACK * ATT_NORM_SLACK);

    // Softmax reference of a unit of this group (CTA unit number itn): |q_r| from the query tile in
    // shared memory, max |k| from the key-norm warp, -m_r into the A slice of the third K-step.  The
    // path (bound or exact maximum) is decided per pair of warps sharing rows; both compute it from the
    // same data.  No barrier: everything a thread needs it reads itself.
    auto reference = [&](uint32_t itn, int4& um_out, bool& bound_out) {
      const int st = itn % ATT_STAGES;
      mbar_wait(&full[st], (itn / ATT_STAGES) & 1);            // tiles and unit record of the unit are visible
      {
        const uint4 m = lds128(s_meta + st * 16);
        um_out = make_int4((int)m.x, (int)m.y, (int)m.z, (int)m.w);
      }
      float q2 = 0.f;
      if (r < q_rows) {
        const uint32_t qr = s_smem + st * ATT_STAGE_BYTES + r * 64;
        const int sw = (r >> 1) & 3;                           // visit the 16-byte chunks in a conflict-free order
        __nv_bfloat162 a0 = __float2bfloat162_rn(0.f), a1 = a0;
#pragma unroll
        for (int g = 0; g < 4; ++g) att_sumsq8(lds128(qr + ((g ^ sw) << 4)), a0, a1);
        q2 = att_sumsq_total(a0, a1);
      }
      const float q2max = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(q2)));
      mbar_wait(&kn_ready[st], (itn / ATT_STAGES) & 1);
      const float k2max = __uint_as_float(lds32(s_kmax2 + st * 4));
      bound_out = !p.exact_max && q2max * k2max <= gap_sq;     // 2 |q| |k| <= ATT_BOUND_GAP for each of the warp's rows
      if (half == 0) {
        const float ms = bound_out ? att_sqrt_fma(q2 * k2max) * ATT_NORM_SLACK : 0.f;
        sts16(s_qx_row, __bfloat16_as_ushort(__float2bfloat16_rn(-ms)));
        fence_proxy_async();
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&qx_ready[grp]);
    };

    uint32_t it = grp, k = 0;                                  // CTA unit number, group unit number
    bool have = (long long)blockIdx.x + (long long)it * gridDim.x < p.n_units;
    int4 um = make_int4(0, 0, 0, 1);
    bool use_bound = true;
    if (have) reference(it, um, use_bound);
    while (have) {
      const int lenh = um.w - cbeg * 32;                       // valid keys counted from this half's first
      const bool all_valid = lenh >= ncl * 32;                 // no padding among this warp's keys (the usual case)
      ATT_STAMP(t1);
      mbar_wait(&s_full[grp], k & 1);
      tc_fence_after();
      ATT_STAMP(t2);
      uint32_t v[2][32];
      float mx = 0.f;
      if (!use_bound) {
        // exact maximum over this half's valid keys, combined with the other half's
        mx = all_valid ? att_pass_max<true>(tS, ncl, lenh, v) : att_pass_max<false>(tS, ncl, lenh, v);
        sts32(s_xm + (half * 128 + r) * 4, __float_as_uint(mx));
        named_bar_sync(pair_bar, 64);
        mx = fmaxf(mx, __uint_as_float(lds32(s_xm + ((half ^ 1) * 128 + r) * 4)));   // finite: key 0 is always valid
      }
      ATT_STAMP(t3);
      // the exponential pass, entered only on this group's turn: group 0 goes first, then they alternate
      if (!p.no_turns) mbar_wait(&xu_turn[grp], grp == 0 ? ((k & 1) ^ 1) : (k & 1));
      ATT_STAMP(t4);
      if (use_bound) {
        if (all_valid) att_pass_exp<true, false, NPOLY>(tS, ncl, lenh, 0.f, v, &xu_turn[grp ^ 1], lane);
        else att_pass_exp<false, false, NPOLY>(tS, ncl, lenh, 0.f, v, &xu_turn[grp ^ 1], lane);
      } else {
        if (all_valid) att_pass_exp<true, true, NPOLY>(tS, ncl, lenh, mx, v, &xu_turn[grp ^ 1], lane);
        else att_pass_exp<false, true, NPOLY>(tS, ncl, lenh, mx, v, &xu_turn[grp ^ 1], lane);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_ready[grp]);
      ATT_STAMP(t5);
      // while the P V product runs: the reference of this group's NEXT unit
      const uint32_t itn = it + 2;
      const bool have_n = (long long)blockIdx.x + (long long)itn * gridDim.x < p.n_units;
      int4 um_n = um;
      bool bound_n = true;
      if (have_n) reference(itn, um_n, bound_n);
      ATT_STAMP(t6);
      // output tile: O / l -> bf16 -> this warp's staging rows -> TMA store
      mbar_wait(&o_full[grp], k & 1);
      tc_fence_after();
      ATT_STAMP(t7);
      uint32_t o[16], lsum;
      tmem_ld16(tO, o);
      tmem_ld1(tL, lsum);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&t_free[grp]);
        tma_store_wait_read();                                 // the previous unit's store has left the staging rows
      }
      __syncwarp();
      {
        const float inv = att_rcp_fma(__uint_as_float(lsum));
#pragma unroll
        for (int g = 0; g < 2; ++g) {
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
      if (lane == 0 && um.z * 128 + wq * 32 < S) {             // whole 32-row blocks are valid or not (S % 32 == 0)
        tma_store_2d(&map_out, out_w, um.y * ATT_DH + half * 16, um.x + um.z * 128 + wq * 32);
        tma_store_commit();
      }
#ifdef B200_ATT_TIMING
      ATT_STAMP(t8);
      ATT_ACC(1, t1, t2); ATT_ACC(2, t2, t3); ATT_ACC(3, t3, t4); ATT_ACC(4, t4, t5); ATT_ACC(0, t5, t6);
      ATT_ACC(5, t6, t7); ATT_ACC(6, t7, t8); ATT_ACC(7, 0, 1);
      if (grp == 0 && lane == 0) atomicAdd(&g_att_timing[8 + w8], (unsigned long long)(t5 - t4));
#endif
      um = um_n;
      use_bound = bound_n;
      it = itn;
      have = have_n;
      ++k;
    }
    if (lane == 0) tma_store_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == W_TMA) {
    tc_fence_after();
    tmem_dealloc(tmem_base, ATT_TMEM_COLS);
  }
}

}  // namespace b200
