// attention_tc.cuh -- K3 on tcgen05: softmax(QK^T/sqrt(dh) + key-padding mask) V for 32-wide heads
// and sequences up to 256 tokens (the MiniLM-L6 shape), one (sequence, head) at a time per CTA.
//
//   scores   S = Q K^T     128 x S tile per MMA pair (M128 x N<=256 x K16, two K steps), fp32 in TMEM
//   softmax  one thread per query row (TMEM lane): row max, then P = 2^((s - max) * scale * log2 e)
//            written back OVER the scores as bf16 (two per 32-bit column) with tcgen05.st
//   output   O = P V       A operand straight from TMEM (no shared-memory round trip for P), V read
//            as an MN-major B operand exactly as TMA lands it ([key][dim] rows of 64 bytes)
//
// Warp roles: 0 TMA producer (Q, K, V of the next (sequence, head) while this one is in softmax),
// 1 MMA issuer, 2 TMEM allocator, 4-7 softmax of query rows 0-127, 8-11 softmax of rows 128-255.
// The tensor work is ~1/8 of the kernel; what bounds it is the exponential (one SFU op per score),
// so everything else is arranged to keep the two softmax warpgroups busy.
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int ATT_DH = 32;
constexpr int ATT_MAX_S = 256;
constexpr int ATT_THREADS = 384;
constexpr int ATT_TILE_BYTES = ATT_MAX_S * ATT_DH * 2;   // 16 KB: one of Q / K / V for one (b, h)
constexpr int ATT_STAGE_BYTES = 3 * ATT_TILE_BYTES;

struct AttParams {
  int S;            // tokens per sequence (multiple of 16, <= 256)
  int H;            // hidden size
  int heads;
  int n_items;      // batch * heads
  const int32_t* lens;
  __nv_bfloat16* ctx;
  int v_mn_major;   // 1: V consumed as landed ([key][dim]); 0: reserved
};

B200_DEVINL float att_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(ATT_THREADS, 1)
attention_tc_kernel(const __grid_constant__ CUtensorMap map_qkv, const AttParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * ATT_STAGE_BYTES);
  uint64_t* full = bars;            // [2] TMA -> MMA: Q, K, V of an item landed
  uint64_t* kv_empty = bars + 2;    // [2] MMA -> TMA: the item's MMAs have read the stage
  uint64_t* s_full = bars + 4;      // [2 row tiles] MMA -> softmax: scores ready
  uint64_t* p_ready = bars + 6;     // [2] softmax -> MMA: probabilities written to TMEM
  uint64_t* o_full = bars + 8;      // [2] MMA -> softmax: output tile ready
  uint64_t* t_free = bars + 10;     // [2] softmax -> MMA: TMEM region of the row tile reusable
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = p.S;
  const int n_mt = (S + 127) >> 7;                       // 128-row query tiles per sequence

  if (warp == 0 && lane == 0) tma_prefetch_desc(&map_qkv);
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&kv_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&p_ready[i], 4);    // one arrival per softmax warp of the row tile
      mbar_init(&o_full[i], 1);
      mbar_init(&t_free[i], 4);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tile_bytes = (uint32_t)S * ATT_DH * 2;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      uint32_t it = 0;
      for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++it) {
        const int st = it & 1, b = item / p.heads, h = item % p.heads;
        mbar_wait(&kv_empty[st], ((it >> 1) & 1) ^ 1);
        uint8_t* base = smem + st * ATT_STAGE_BYTES;
        mbar_arrive_expect_tx(&full[st], 3 * tile_bytes);
        tma_load_2d(base, &map_qkv, &full[st], h * ATT_DH, b * S);                                   // Q
        tma_load_2d(base + ATT_TILE_BYTES, &map_qkv, &full[st], p.H + h * ATT_DH, b * S);            // K
        tma_load_2d(base + 2 * ATT_TILE_BYTES, &map_qkv, &full[st], 2 * p.H + h * ATT_DH, b * S);    // V
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      const uint32_t idesc_s = umma_idesc_bf16_f32(128, (uint32_t)S);        // scores: N = S keys
      const uint32_t idesc_o = umma_idesc_bf16_f32_bmn(128, ATT_DH);         // output: N = 32 dims, V MN-major
      uint32_t it = 0;
      for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++it) {
        const int st = it & 1;
        const uint32_t sbase = smem_u32(smem + st * ATT_STAGE_BYTES);
        mbar_wait(&full[st], (it >> 1) & 1);
        tc_fence_after();
        // K as a K-major B operand: rows (keys) of 64 bytes, 8-row swizzle atoms of 512 bytes
        for (int mt = 0; mt < n_mt; ++mt) {
          mbar_wait(&t_free[mt], (it & 1) ^ 1);          // the previous item's tile mt has been drained
          tc_fence_after();
          const uint32_t tS = tmem_base + mt * 256;
#pragma unroll
          for (int k = 0; k < ATT_DH / 16; ++k) {
            const uint64_t dq = umma_smem_desc(sbase + (uint32_t)mt * 128 * 64 + k * 32, 4, 512, 0);
            const uint64_t dk = umma_smem_desc(sbase + ATT_TILE_BYTES + k * 32, 4, 512, 0);
            umma_bf16(tS, dq, dk, idesc_s, (uint32_t)(k != 0));
          }
          umma_commit(&s_full[mt]);
        }
        for (int mt = 0; mt < n_mt; ++mt) {
          mbar_wait(&p_ready[mt], it & 1);
          tc_fence_after();
          const uint32_t tP = tmem_base + mt * 256;             // bf16 pairs in columns [0, S/2)
          const uint32_t tO = tmem_base + mt * 256 + 128;       // fp32 output, 32 columns
          const int ksteps = S / 16;
          for (int k = 0; k < ksteps; ++k) {
            // V as an MN-major B operand: 16 keys per step = 16 rows of 64 bytes = 1024 bytes
            const uint64_t dv = umma_smem_desc(sbase + 2 * ATT_TILE_BYTES + k * 1024, 4, 512, 16384);
            umma_bf16_ts(tO, tP + k * 8, dv, idesc_o, (uint32_t)(k != 0));
          }
          umma_commit(&o_full[mt]);
        }
        umma_commit(&kv_empty[st]);   // every MMA that reads this stage has been issued; free it when they finish
      }
    }
  } else if (warp >= 4) {
    // ===== softmax + output: thread = one query row of row tile mt =====
    const int mt = (warp - 4) >> 2;
    const int wq = warp & 3;                                   // TMEM lane group of this warp
    if (mt < n_mt) {
      const int row = mt * 128 + wq * 32 + lane;               // query position in the sequence
      const uint32_t tS = tmem_base + ((uint32_t)(wq * 32) << 16) + mt * 256;
      const float scale_log2 = rsqrtf((float)ATT_DH) * 1.4426950408889634f;
      const int nch = S / 32;                                  // 32-column chunks (S multiple of 32 here)
      uint32_t it = 0;
      for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++it) {
        const int b = item / p.heads, h = item % p.heads;
        const int len = max(1, min(p.lens[b], S));
        mbar_wait(&s_full[mt], it & 1);
        tc_fence_after();
        // pass 1: row maximum over the valid keys
        float mx = -INFINITY;
        for (int c = 0; c < nch; ++c) {
          if (c * 32 >= len) break;
          uint32_t v[32];
          tmem_ld32(tS + c * 32, v);
          tmem_ld_wait();
          if (c * 32 + 32 <= len) {
#pragma unroll
            for (int j = 0; j < 32; ++j) mx = fmaxf(mx, __uint_as_float(v[j]));
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) mx = fmaxf(mx, (c * 32 + j < len) ? __uint_as_float(v[j]) : -INFINITY);
          }
        }
        const float ms = mx * scale_log2;
        // pass 2: probabilities (bf16, packed) written back over the scores; row sum in fp32
        float sum = 0.f;
        for (int c = 0; c < nch; ++c) {
          uint32_t pk[16];
          if (c * 32 < len) {
            uint32_t v[32];
            tmem_ld32(tS + c * 32, v);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              float p0 = att_exp2(fmaf(__uint_as_float(v[2 * j]), scale_log2, -ms));
              float p1 = att_exp2(fmaf(__uint_as_float(v[2 * j + 1]), scale_log2, -ms));
              if (c * 32 + 32 > len) {
                if (c * 32 + 2 * j >= len) p0 = 0.f;
                if (c * 32 + 2 * j + 1 >= len) p1 = 0.f;
              }
              sum += p0 + p1;
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
        if (lane == 0) mbar_arrive(&p_ready[mt]);
        // output tile: O / sum -> bf16 -> ctx[b, row, h*32 .. +32]
        mbar_wait(&o_full[mt], it & 1);
        tc_fence_after();
        uint32_t o[32];
        tmem_ld32(tS + 128, o);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&t_free[mt]);
        if (row < S) {
          const float inv = 1.0f / sum;
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
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace b200
