// attention_tc.cuh -- K3 on tcgen05: softmax(QK^T/sqrt(dh) + key-padding mask) V for 32-wide heads
// and sequences up to 256 tokens (the MiniLM-L6 shape).
//
// Work unit = (sequence, head, 128-row query tile).  Per unit:
//   scores   S = Q K^T     one 128 x S tile (M128 x N<=256 x K16, two K steps), fp32 in TMEM
//   softmax  one thread per query row (TMEM lane): row max, then P = 2^((s - max) * scale * log2 e)
//            written back OVER the scores as bf16 (two per 32-bit column) with tcgen05.st
//   output   O = P V       A operand straight from TMEM (no shared-memory round trip for P), V read
//            as an MN-major B operand exactly as TMA lands it ([key][dim] rows of 64 bytes)
//
// Warp roles: 0 TMA producer (Q, K, V of the next unit while this one is in softmax), 1 MMA issuer,
// 2 TMEM allocator, 4-7 softmax + output.  A CTA uses 256 TMEM columns and ~82 KB of shared memory, so
// TWO CTAs share an SM: while one waits on its MMA round trips (scores ready, output ready) the other
// one's softmax warps keep the issue slots and the SFU busy.  The tensor work is ~1/8 of the kernel;
// what bounds it is the exponential (one SFU op per score).
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int ATT_DH = 32;
constexpr int ATT_MAX_S = 256;
constexpr int ATT_THREADS = 256;
constexpr int ATT_Q_BYTES = 128 * ATT_DH * 2;            // 8 KB: one 128-row query tile
constexpr int ATT_KV_BYTES = ATT_MAX_S * ATT_DH * 2;     // 16 KB: K or V of one (sequence, head)
constexpr int ATT_STAGE_BYTES = ATT_Q_BYTES + 2 * ATT_KV_BYTES;
constexpr int ATT_TMEM_COLS = 256;

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
__device__ unsigned long long g_att_timing[8];
#define ATT_STAMP(var) const long long var = clock64()
#define ATT_ACC(i, a, b) do { if (wq == 0 && lane == 0) atomicAdd(&g_att_timing[i], (unsigned long long)((b) - (a))); } while (0)
#else
#define ATT_STAMP(var)
#define ATT_ACC(i, a, b)
#endif

B200_DEVINL float att_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Probabilities of one 32-key chunk of a row: p = 2^(s*scale - ms) for the first `valid` keys, 0 for
// the rest (padding); accumulates the row sum and packs bf16 pairs.
B200_DEVINL void att_probs(const uint32_t (&v)[32], int valid, float scale_log2, float ms, float& sum0, float& sum1,
                           uint32_t (&pk)[16]) {
  if (valid >= 32) {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const float p0 = att_exp2(fmaf(__uint_as_float(v[2 * j]), scale_log2, -ms));
      const float p1 = att_exp2(fmaf(__uint_as_float(v[2 * j + 1]), scale_log2, -ms));
      sum0 += p0;
      sum1 += p1;
      __nv_bfloat162 t = __floats2bfloat162_rn(p0, p1);
      pk[j] = *reinterpret_cast<uint32_t*>(&t);
    }
  } else if (valid > 0) {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      float p0 = att_exp2(fmaf(__uint_as_float(v[2 * j]), scale_log2, -ms));
      float p1 = att_exp2(fmaf(__uint_as_float(v[2 * j + 1]), scale_log2, -ms));
      if (2 * j >= valid) p0 = 0.f;
      if (2 * j + 1 >= valid) p1 = 0.f;
      sum0 += p0;
      sum1 += p1;
      __nv_bfloat162 t = __floats2bfloat162_rn(p0, p1);
      pk[j] = *reinterpret_cast<uint32_t*>(&t);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 16; ++j) pk[j] = 0u;
  }
}

__global__ void __launch_bounds__(ATT_THREADS, 2)
attention_tc_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_kv,
                    const AttParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * ATT_STAGE_BYTES);
  uint64_t* full = bars;            // [2] TMA -> MMA: Q, K, V of a unit landed
  uint64_t* kv_empty = bars + 2;    // [2] MMA -> TMA: the unit's MMAs have read the stage
  uint64_t* s_full = bars + 4;      // MMA -> softmax: scores ready
  uint64_t* o_full = bars + 5;      // MMA -> softmax: output tile ready
  uint64_t* t_free = bars + 6;      // softmax -> MMA: TMEM reusable
  uint64_t* p_ready = bars + 8;     // [4] softmax -> MMA: probabilities of one 64-key group written to TMEM
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);
  int4* meta = reinterpret_cast<int4*>(bars + 14);   // [2] per stage: {token row base, head, query tile, length}

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = p.S;
  const int q_rows = min(S, 128);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_q);
    tma_prefetch_desc(&map_kv);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    mbar_init(s_full, 1);
    for (int i = 0; i < 4; ++i) mbar_init(&p_ready[i], 4);    // one arrival per softmax warp
    mbar_init(o_full, 1);
    mbar_init(t_free, 4);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, ATT_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t kv_bytes = (uint32_t)S * ATT_DH * 2, q_bytes = (uint32_t)q_rows * ATT_DH * 2;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      uint32_t it = 0;
      for (int u = blockIdx.x; u < p.n_units; u += gridDim.x, ++it) {
        const int st = it & 1, item = u / p.n_mt, mt = u - item * p.n_mt;
        const int b = item / p.heads, h = item - b * p.heads;
        const int len = p.lens[b];
        mbar_wait(&kv_empty[st], ((it >> 1) & 1) ^ 1);
        uint8_t* base = smem + st * ATT_STAGE_BYTES;
        meta[st] = make_int4(b * S, h, mt, max(1, min(len, S)));   // published by the arrive below
        mbar_arrive_expect_tx(&full[st], q_bytes + 2 * kv_bytes);
        tma_load_2d(base, &map_q, &full[st], h * ATT_DH, b * S + mt * 128);                                  // Q tile
        tma_load_2d(base + ATT_Q_BYTES, &map_kv, &full[st], p.H + h * ATT_DH, b * S);                        // K
        tma_load_2d(base + ATT_Q_BYTES + ATT_KV_BYTES, &map_kv, &full[st], 2 * p.H + h * ATT_DH, b * S);     // V
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      const uint32_t idesc_s = umma_idesc_bf16_f32(128, (uint32_t)S);        // scores: N = S keys
      const uint32_t idesc_o = umma_idesc_bf16_f32_bmn(128, ATT_DH);         // output: N = 32 dims, V MN-major
      const uint32_t tS = tmem_base;                 // scores fp32, columns [0, S)
      const uint32_t tO = tmem_base;                 // output fp32, columns [0, 32) once scores chunk 0 is consumed
      const uint32_t tP = tmem_base + 32;            // probabilities, bf16 pairs: 8 columns per 16-key step
      uint32_t it = 0;
      for (int u = blockIdx.x; u < p.n_units; u += gridDim.x, ++it) {
        const int st = it & 1;
        const uint32_t sbase = smem_u32(smem + st * ATT_STAGE_BYTES);
        mbar_wait(&full[st], (it >> 1) & 1);
        mbar_wait(t_free, (it & 1) ^ 1);             // the previous unit's output has been drained
        tc_fence_after();
        // Q (A) and K (B) are K-major: rows of 64 bytes, 8-row swizzle atoms of 512 bytes
#pragma unroll
        for (int k = 0; k < ATT_DH / 16; ++k) {
          const uint64_t dq = umma_smem_desc(sbase + k * 32, 4, 512, 0);
          const uint64_t dk = umma_smem_desc(sbase + ATT_Q_BYTES + k * 32, 4, 512, 0);
          umma_bf16(tS, dq, dk, idesc_s, (uint32_t)(k != 0));
        }
        umma_commit(s_full);
        // P V in groups of 64 keys, each issued as soon as the softmax warps have written that group's
        // probabilities, so that only the last group's MMAs remain when the softmax pass ends
        const int ksteps = S / 16;
        for (int g = 0; g * 4 < ksteps; ++g) {
          mbar_wait(&p_ready[g], it & 1);
          tc_fence_after();
          const int k1 = min(ksteps, g * 4 + 4);
          for (int k = g * 4; k < k1; ++k) {
            // V as an MN-major B operand: 16 keys per step = 16 rows of 64 bytes = 1024 bytes
            const uint64_t dv = umma_smem_desc(sbase + ATT_Q_BYTES + ATT_KV_BYTES + k * 1024, 4, 512, 16384);
            umma_bf16_ts(tO, tP + k * 8, dv, idesc_o, (uint32_t)(k != 0));
          }
        }
        umma_commit(o_full);
        umma_commit(&kv_empty[st]);   // the stage is free once every MMA issued so far has finished
      }
    }
  } else if (warp >= 4) {
    // ===== softmax + output: thread = one query row =====
    const int wq = warp & 3;                                   // TMEM lane group of this warp
    const uint32_t tS = tmem_base + ((uint32_t)(wq * 32) << 16);
    const float scale_log2 = rsqrtf((float)ATT_DH) * 1.4426950408889634f;
    const int nch = S / 32;
    uint32_t it = 0;
    for (int u = blockIdx.x; u < p.n_units; u += gridDim.x, ++it) {
      const int st = it & 1;
      ATT_STAMP(t0);
      mbar_wait(&full[st], (it >> 1) & 1);                     // the producer's unit record is visible
      const int4 um = meta[st];
      const int row = um.z * 128 + wq * 32 + lane;             // query position in the sequence
      const int len = um.w;
      __nv_bfloat16* dst = p.ctx + ((size_t)um.x + row) * p.H + um.y * ATT_DH;
      mbar_wait(s_full, it & 1);
      tc_fence_after();
      ATT_STAMP(t1);
      // Both passes walk the row in 32-column chunks with the tcgen05.ld of chunk c+1 in flight while
      // chunk c is processed (two register buffers); keys >= len only ever sit in the last chunks.
      const int nfull = len >> 5;                      // chunks made of valid keys only
      const int rem = len & 31;                        // valid keys in the boundary chunk (0 = none)
      uint32_t v[2][32];
      // pass 1: row maximum over the valid keys
      float mx = -INFINITY;
      tmem_ld32(tS, v[0]);
#pragma unroll 1
      for (int c2 = 0; c2 < nfull; c2 += 2) {
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int c = c2 + u;
          if (c < nfull) {
            tmem_ld_wait_dep(v[u]);
            if (c + 1 < nch) tmem_ld32(tS + (c + 1) * 32, v[u ^ 1]);
            float m0 = __uint_as_float(v[u][0]), m1 = __uint_as_float(v[u][1]);
#pragma unroll
            for (int j = 2; j < 32; j += 2) {
              m0 = fmaxf(m0, __uint_as_float(v[u][j]));
              m1 = fmaxf(m1, __uint_as_float(v[u][j + 1]));
            }
            mx = fmaxf(mx, fmaxf(m0, m1));
          }
        }
      }
      if (rem) {                                       // boundary chunk nfull sits in buffer nfull & 1
        if (nfull & 1) {
          tmem_ld_wait_dep(v[1]);
#pragma unroll
          for (int j = 0; j < 32; ++j) mx = fmaxf(mx, (j < rem) ? __uint_as_float(v[1][j]) : -INFINITY);
        } else {
          tmem_ld_wait_dep(v[0]);
#pragma unroll
          for (int j = 0; j < 32; ++j) mx = fmaxf(mx, (j < rem) ? __uint_as_float(v[0][j]) : -INFINITY);
        }
      } else {
        tmem_ld_wait();                                // drain a load that may still be in flight
      }
      const float ms = mx * scale_log2;
      ATT_STAMP(t2);
      // pass 2: probabilities (bf16, packed two per column) and the fp32 row sum.  TMEM layout from
      // here on: output accumulator in columns [0, 32) (scores chunk 0, consumed first), probabilities
      // of chunk c in columns [32 + 16c, 48 + 16c) -- always below the scores still unread, except for
      // chunk 0, whose store therefore waits until chunk 1 has been drained into registers.
      float sum0 = 0.f, sum1 = 0.f;
      const uint32_t tP = tS + 32;
      tmem_ld32(tS, v[0]);
#pragma unroll 1
      for (int c = 0, g = 0; c < nch; c += 2, ++g) {
        uint32_t pk0[16], pk1[16];
        tmem_ld_wait_dep(v[0]);
        if (c + 1 < nch) tmem_ld32(tS + (c + 1) * 32, v[1]);
        att_probs(v[0], len - c * 32, scale_log2, ms, sum0, sum1, pk0);
        if (c + 1 < nch) {
          tmem_ld_wait_dep(v[1]);
          tmem_st16(tP + c * 16, pk0);
          if (c + 2 < nch) tmem_ld32(tS + (c + 2) * 32, v[0]);
          att_probs(v[1], len - (c + 1) * 32, scale_log2, ms, sum0, sum1, pk1);
          tmem_st16(tP + (c + 1) * 16, pk1);
        } else {
          tmem_st16(tP + c * 16, pk0);
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_ready[g]);     // this 64-key group may enter the P V product
      }
      // output tile: O / sum -> bf16 -> ctx[b, row, h*32 .. +32]
      ATT_STAMP(t3);
      mbar_wait(o_full, it & 1);
      tc_fence_after();
      ATT_STAMP(t4);
      uint32_t o[32];
      tmem_ld32(tS, o);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(t_free);
      if (row < S) {
        const float inv = 1.0f / (sum0 + sum1);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 w;
          __nv_bfloat162 t0 = __floats2bfloat162_rn(__uint_as_float(o[g * 8 + 0]) * inv, __uint_as_float(o[g * 8 + 1]) * inv);
          __nv_bfloat162 t1 = __floats2bfloat162_rn(__uint_as_float(o[g * 8 + 2]) * inv, __uint_as_float(o[g * 8 + 3]) * inv);
          __nv_bfloat162 t2 = __floats2bfloat162_rn(__uint_as_float(o[g * 8 + 4]) * inv, __uint_as_float(o[g * 8 + 5]) * inv);
          __nv_bfloat162 t3 = __floats2bfloat162_rn(__uint_as_float(o[g * 8 + 6]) * inv, __uint_as_float(o[g * 8 + 7]) * inv);
          w.x = *reinterpret_cast<uint32_t*>(&t0); w.y = *reinterpret_cast<uint32_t*>(&t1);
          w.z = *reinterpret_cast<uint32_t*>(&t2); w.w = *reinterpret_cast<uint32_t*>(&t3);
          *reinterpret_cast<uint4*>(dst + g * 8) = w;
        }
      }
#ifdef B200_ATT_TIMING
      ATT_STAMP(t5);
      ATT_ACC(0, t0, t1); ATT_ACC(1, t1, t2); ATT_ACC(2, t2, t3); ATT_ACC(3, t3, t4); ATT_ACC(4, t4, t5); ATT_ACC(5, 0, 1);
#endif
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, ATT_TMEM_COLS);
  }
}

}  // namespace b200
