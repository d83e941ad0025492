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

B200_DEVINL float att_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
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
  uint64_t* p_ready = bars + 5;     // softmax -> MMA: probabilities written to TMEM
  uint64_t* o_full = bars + 6;      // MMA -> softmax: output tile ready
  uint64_t* t_free = bars + 7;      // softmax -> MMA: TMEM reusable
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

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
    mbar_init(p_ready, 4);    // one arrival per softmax warp
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
        mbar_wait(&kv_empty[st], ((it >> 1) & 1) ^ 1);
        uint8_t* base = smem + st * ATT_STAGE_BYTES;
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
      const uint32_t tS = tmem_base;                 // scores fp32 [0, S); later P bf16 pairs [0, S/2)
      const uint32_t tO = tmem_base + 128;           // output fp32, 32 columns
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
        mbar_wait(p_ready, it & 1);
        tc_fence_after();
        const int ksteps = S / 16;
        for (int k = 0; k < ksteps; ++k) {
          // V as an MN-major B operand: 16 keys per step = 16 rows of 64 bytes = 1024 bytes
          const uint64_t dv = umma_smem_desc(sbase + ATT_Q_BYTES + ATT_KV_BYTES + k * 1024, 4, 512, 16384);
          umma_bf16_ts(tO, tS + k * 8, dv, idesc_o, (uint32_t)(k != 0));
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
    int len_next = 1;
    if ((int)blockIdx.x < p.n_units) len_next = p.lens[(blockIdx.x / p.n_mt) / p.heads];
    for (int u = blockIdx.x; u < p.n_units; u += gridDim.x, ++it) {
      const int item = u / p.n_mt, mt = u - item * p.n_mt;
      const int b = item / p.heads, h = item - b * p.heads;
      const int row = mt * 128 + wq * 32 + lane;               // query position in the sequence
      const int len = max(1, min(len_next, S));
      if (u + (int)gridDim.x < p.n_units) len_next = p.lens[((u + gridDim.x) / p.n_mt) / p.heads];   // hide the load
      mbar_wait(s_full, it & 1);
      tc_fence_after();
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
      // pass 2: probabilities (bf16, packed) written back over the scores; row sum in fp32
      float sum0 = 0.f, sum1 = 0.f;
      tmem_ld32(tS, v[0]);
#pragma unroll 1
      for (int c2 = 0; c2 < nfull; c2 += 2) {
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int c = c2 + u;
          if (c < nfull) {
            tmem_ld_wait_dep(v[u]);
            if (c + 1 < nch) tmem_ld32(tS + (c + 1) * 32, v[u ^ 1]);
            uint32_t pk[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float p0 = att_exp2(fmaf(__uint_as_float(v[u][2 * j]), scale_log2, -ms));
              const float p1 = att_exp2(fmaf(__uint_as_float(v[u][2 * j + 1]), scale_log2, -ms));
              sum0 += p0;
              sum1 += p1;
              __nv_bfloat162 t = __floats2bfloat162_rn(p0, p1);
              pk[j] = *reinterpret_cast<uint32_t*>(&t);
            }
            tmem_st16(tS + c * 16, pk);
          }
        }
      }
      for (int c = nfull; c < nch; ++c) {              // boundary chunk (masked), then all-padding chunks (zeros)
        uint32_t pk[16];
        if (c == nfull && rem) {
          uint32_t w[32];
          tmem_ld32(tS + c * 32, w);                   // (re-)load: keeps this rare path simple
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            float p0 = att_exp2(fmaf(__uint_as_float(w[2 * j]), scale_log2, -ms));
            float p1 = att_exp2(fmaf(__uint_as_float(w[2 * j + 1]), scale_log2, -ms));
            if (2 * j >= rem) p0 = 0.f;
            if (2 * j + 1 >= rem) p1 = 0.f;
            sum0 += p0;
            sum1 += p1;
            __nv_bfloat162 t = __floats2bfloat162_rn(p0, p1);
            pk[j] = *reinterpret_cast<uint32_t*>(&t);
          }
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) pk[j] = 0u;
        }
        tmem_st16(tS + c * 16, pk);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_ready);
      // output tile: O / sum -> bf16 -> ctx[b, row, h*32 .. +32]
      mbar_wait(o_full, it & 1);
      tc_fence_after();
      uint32_t o[32];
      tmem_ld32(tS + 128, o);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(t_free);
      if (row < S) {
        const float inv = 1.0f / (sum0 + sum1);
        __nv_bfloat16* dst = p.ctx + ((size_t)b * S + row) * p.H + h * ATT_DH;
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
