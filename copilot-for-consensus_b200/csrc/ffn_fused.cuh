// ffn_fused.cuh -- K5 + K6 in one kernel for hidden 384 (MiniLM-class):
//   x <- LayerNorm( x + gelu(x W1^T + b1) W2^T + b2 )
// without the [tokens, intermediate] activation ever leaving the SM.  As two kernels the intermediate
// (403 MB at 131072 tokens) is written by one, evicted from the 126 MB L2, and read back by the other;
// the FFN1 store path alone costs ~60 us per layer (measured by switching the stores off).
//
// A CTA pair (cta_group::2) owns a 256-row block and walks the intermediate in slices of 128 columns:
//   G1(j)  H  = X W1[j]^T          M256 x N128 x K384   A: the block's rows of x, resident in shared memory
//   E1(j)  Hs = bf16(gelu(H + b1))  epilogue warps: TMEM -> registers -> shared memory, written in the
//                                   128-byte-swizzled K-major layout a TMA load would have produced
//   G2(j)  Y += Hs W2[:, j]^T      M256 x N384 (two N = 192 halves) x K128, A: Hs
// and finishes with the residual + LayerNorm epilogue of gemm_ln_kernel over the 384 accumulator columns.
// The MMA issuer runs G1(j+1) before G2(j), so the tensor pipe works on the next slice while the
// epilogue warps turn slice j around; TMEM = 384 columns of Y + 128 of H = all 512.
//
// Warp roles: 0 TMA producer (x slab k-blocks once per row block, W1 / W2 tiles through two rings),
// 1 MMA issuer (leader CTA), 2 TMEM allocator, 4-11 epilogue (warp e: TMEM lane group e & 3, column
// half e >> 2).  The 4 KB of the Hs tile that warp e writes in E1 is also its staging box for the
// residual load and the output store of the final epilogue, so no other warp ever touches it.
#pragma once
#include "encoder_kernels.cuh"

namespace b200 {

constexpr int FFN_H = 384;                                // hidden size this kernel is written for
constexpr int FFN_SLICE = 128;                            // intermediate columns per slice
constexpr int FFN_KB1 = FFN_H / GEMM_BLOCK_K;             // 6 k-blocks of G1
constexpr int FFN_KB2 = FFN_SLICE / GEMM_BLOCK_K;         // 2 k-blocks of G2 per slice
#ifndef FFN_W1_NSTAGE
#define FFN_W1_NSTAGE 5
#endif
constexpr int FFN_W1_STAGES = FFN_W1_NSTAGE;
constexpr int FFN_W2_STAGES = 2;
constexpr int FFN_EPI_WARPS = 8;
constexpr int FFN_THREADS = (4 + FFN_EPI_WARPS) * 32;
constexpr uint32_t FFN_X_BYTES = FFN_KB1 * GEMM_A_STAGE_BYTES;             // 96 KB: 128 rows x 384
constexpr uint32_t FFN_W1_STAGE_BYTES = (FFN_SLICE / 2) * GEMM_BLOCK_K * 2;  // 8 KB: this CTA's 64 rows of a W1 k-block
constexpr uint32_t FFN_W2_HALF_BYTES = (192 / 2) * GEMM_BLOCK_K * 2;         // 12 KB: this CTA's 96 rows of one N half
constexpr uint32_t FFN_W2_STAGE_BYTES = 2 * FFN_W2_HALF_BYTES;               // 24 KB: both N halves of a W2 k-block
constexpr uint32_t FFN_HS_BYTES = FFN_KB2 * GEMM_A_STAGE_BYTES;              // 32 KB: 128 rows x 128 bf16
constexpr uint32_t FFN_OFF_W1 = FFN_X_BYTES;
constexpr uint32_t FFN_OFF_W2 = FFN_OFF_W1 + FFN_W1_STAGES * FFN_W1_STAGE_BYTES;
constexpr uint32_t FFN_OFF_HS = FFN_OFF_W2 + FFN_W2_STAGES * FFN_W2_STAGE_BYTES;
constexpr uint32_t FFN_OFF_TAIL = FFN_OFF_HS + FFN_HS_BYTES;
// tail: barriers (64 x 8 B), TMEM slot, b2 / gamma / beta [3][384], row statistics [2 parities][2][128]{sum, sumsq}
constexpr uint32_t FFN_TAIL_BYTES = 512 + 16 + 3 * FFN_H * 4 + 2 * 2 * 128 * 8;
constexpr size_t FFN_SMEM_BYTES = 1024 + FFN_OFF_TAIL + FFN_TAIL_BYTES;

struct FfnParams {
  int M;               // rows (multiple of 256)
  int I;               // intermediate size (multiple of FFN_SLICE)
  const float* b1;     // [I]
  const float* b2;     // [384]
  const float* gamma;  // [384]
  const float* beta;   // [384]
  float eps;
  int reverse;         // see GemmParams::reverse
};

__global__ void __launch_bounds__(FFN_THREADS, 1)
ffn_fused_kernel(const __grid_constant__ CUtensorMap map_x,      // x as A operand: box {64 cols, 128 rows}
                 const __grid_constant__ CUtensorMap map_w1,     // W1 [I, 384]: box {64, 64}
                 const __grid_constant__ CUtensorMap map_w2,     // W2 [384, I]: box {64, 96}
                 const __grid_constant__ CUtensorMap map_xio,    // x as residual in / output: box {64, 32}
                 const FfnParams p) {
  constexpr int CG = 2;
  constexpr uint32_t IDESC1 = umma_idesc_bf16_f32(GEMM_BLOCK_M * CG, FFN_SLICE);
  constexpr uint32_t IDESC2 = umma_idesc_bf16_f32(GEMM_BLOCK_M * CG, 192);
  constexpr uint32_t TMEM_Y = 0, TMEM_H = FFN_H;           // columns: 384 of Y, then 128 of H

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* w1_ring = smem + FFN_OFF_W1;
  uint8_t* w2_ring = smem + FFN_OFF_W2;
  uint8_t* hs = smem + FFN_OFF_HS;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + FFN_OFF_TAIL);
  uint64_t* x_full = bars;                    // [6]  TMA -> MMA   (leader's copy)     slab k-block landed
  uint64_t* x_empty = bars + 6;               // [6]  MMA -> TMA   (one per CTA)       last G1 of the row block read it
  uint64_t* w1_full = bars + 32;              // [FFN_W1_STAGES <= 8]  TMA -> MMA   (leader's copy)
  uint64_t* w1_empty = bars + 40;             // [FFN_W1_STAGES <= 8]  MMA -> TMA   (one per CTA)
  uint64_t* w2_full = bars + 18;              // [2]  TMA -> MMA   (leader's copy)
  uint64_t* w2_empty = bars + 20;             // [2]  MMA -> TMA   (one per CTA)
  uint64_t* h_full = bars + 22;               // MMA -> epilogue   (one per CTA)       G1(j) complete
  uint64_t* h_free = bars + 23;               // epilogue -> MMA   (leader's copy)     H accumulator drained
  uint64_t* hs_full = bars + 24;              // epilogue -> MMA   (leader's copy)     Hs written
  uint64_t* hs_free = bars + 25;              // MMA -> epilogue   (one per CTA)       G2(j) has read Hs
  uint64_t* y_full = bars + 26;               // MMA -> epilogue   (one per CTA)       row block's Y complete
  uint64_t* y_free = bars + 27;               // epilogue -> MMA   (leader's copy)     Y drained
  uint64_t* resid_bar = bars + 48;            // [8]  residual box landed (one per epilogue warp)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 64);
  float* vecs = reinterpret_cast<float*>(tmem_slot + 4);                   // b2, gamma, beta: [3][384]
  float* xs = vecs + 3 * FFN_H;                                            // [2 parities][2][128] {sum, sumsq}

  const int warp = uniform_warp_idx();
  const int lane = threadIdx.x & 31;
  const int rank = (int)cluster_ctarank();
  const bool leader = rank == 0;
  const int m_blocks = p.M / (GEMM_BLOCK_M * CG);
  const int group = blockIdx.x / CG, n_groups = gridDim.x / CG;
  const int n_slices = p.I / FFN_SLICE;

  for (int i = threadIdx.x; i < FFN_H; i += blockDim.x) {
    vecs[i] = __ldg(p.b2 + i);
    vecs[FFN_H + i] = __ldg(p.gamma + i);
    vecs[2 * FFN_H + i] = __ldg(p.beta + i);
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_x);
    tma_prefetch_desc(&map_w1);
    tma_prefetch_desc(&map_w2);
    tma_prefetch_desc(&map_xio);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < FFN_KB1; ++i) { mbar_init(&x_full[i], CG); mbar_init(&x_empty[i], 1); }
    for (int i = 0; i < FFN_W1_STAGES; ++i) { mbar_init(&w1_full[i], CG); mbar_init(&w1_empty[i], 1); }
    for (int i = 0; i < FFN_W2_STAGES; ++i) { mbar_init(&w2_full[i], CG); mbar_init(&w2_empty[i], 1); }
    mbar_init(h_full, 1);
    mbar_init(h_free, FFN_EPI_WARPS * CG);
    mbar_init(hs_full, FFN_EPI_WARPS * CG);
    mbar_init(hs_free, 1);
    mbar_init(y_full, 1);
    mbar_init(y_free, FFN_EPI_WARPS * CG);
    for (int i = 0; i < FFN_EPI_WARPS; ++i) mbar_init(&resid_bar[i], 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc_2sm(tmem_slot, 512);
  tc_fence_before();
  cluster_sync_all();                          // the peer's barriers are initialised before anyone signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();
  pdl_launch_dependents();

  if (warp == 0) {
    // ===== TMA producer (converged loop, one elected lane issues).  Load order = the MMA issuer's
    // consumption order: W1 tiles of slice j (with the x slab when j == 0), then W2 tiles of slice j-1. =====
    int s1 = 0, s2 = 0;
    uint32_t ph1 = 0, ph2 = 0, mbi = 0;
    auto load_w2 = [&](int jj) {
      for (int kb2 = 0; kb2 < FFN_KB2; ++kb2) {
        mbar_wait(&w2_empty[s2], ph2 ^ 1);
        if (elect_one()) {
          if (leader) mbar_arrive_expect_tx(&w2_full[s2], FFN_W2_STAGE_BYTES * CG);
          else mbar_arrive_leader(&w2_full[s2]);
          uint8_t* dst = w2_ring + (size_t)s2 * FFN_W2_STAGE_BYTES;
          const int kcol = jj * FFN_SLICE + kb2 * GEMM_BLOCK_K;
          tma_load_2d_2sm(dst, &map_w2, &w2_full[s2], kcol, rank * 96);                              // N half 0
          tma_load_2d_2sm(dst + FFN_W2_HALF_BYTES, &map_w2, &w2_full[s2], kcol, 192 + rank * 96);    // N half 1
        }
        __syncwarp();
        if (++s2 == FFN_W2_STAGES) { s2 = 0; ph2 ^= 1; }
      }
    };
    for (int mb = group; mb < m_blocks; mb += n_groups, ++mbi) {
      const int m0 = (p.reverse ? m_blocks - 1 - mb : mb) * (GEMM_BLOCK_M * CG) + rank * GEMM_BLOCK_M;
      for (int j = 0; j < n_slices; ++j) {
        for (int kb = 0; kb < FFN_KB1; ++kb) {
          if (j == 0) {
            mbar_wait(&x_empty[kb], (mbi & 1) ^ 1);        // the previous row block is done with this k-block
            if (elect_one()) {
              if (leader) mbar_arrive_expect_tx(&x_full[kb], GEMM_A_STAGE_BYTES * CG);
              else mbar_arrive_leader(&x_full[kb]);
              tma_load_2d_2sm(smem + (size_t)kb * GEMM_A_STAGE_BYTES, &map_x, &x_full[kb], kb * GEMM_BLOCK_K, m0);
            }
            __syncwarp();
          }
          mbar_wait(&w1_empty[s1], ph1 ^ 1);
          if (elect_one()) {
            if (leader) mbar_arrive_expect_tx(&w1_full[s1], FFN_W1_STAGE_BYTES * CG);
            else mbar_arrive_leader(&w1_full[s1]);
            tma_load_2d_2sm(w1_ring + (size_t)s1 * FFN_W1_STAGE_BYTES, &map_w1, &w1_full[s1], kb * GEMM_BLOCK_K,
                            j * FFN_SLICE + rank * (FFN_SLICE / 2));
          }
          __syncwarp();
          if (++s1 == FFN_W1_STAGES) { s1 = 0; ph1 ^= 1; }
        }
        if (j > 0) load_w2(j - 1);
      }
      load_w2(n_slices - 1);
    }
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA; converged loop, one elected lane issues): G1(j), then G2(j-1) =====
    if (leader) {
      const uint32_t x_addr = smem_u32(smem), w1_addr = smem_u32(w1_ring), w2_addr = smem_u32(w2_ring), hs_addr = smem_u32(hs);
      int s1 = 0, s2 = 0;
      uint32_t ph1 = 0, ph2 = 0, mbi = 0;
      uint32_t g1c = 0, g2c = 0;              // slices issued so far (all row blocks): phases of the H / Hs hand-shakes
      auto issue_g2 = [&](int jj) {
        mbar_wait(hs_full, g2c & 1);                       // the epilogue warps have written Hs of this slice
        if (jj == 0) mbar_wait(y_free, (mbi & 1) ^ 1);     // the previous row block's Y has been drained
        tc_fence_after();
        for (int kb2 = 0; kb2 < FFN_KB2; ++kb2) {
          mbar_wait(&w2_full[s2], ph2);
          tc_fence_after();
          const uint64_t da = umma_smem_desc_sw128(hs_addr + (uint32_t)kb2 * GEMM_A_STAGE_BYTES);
          const uint64_t db0 = umma_smem_desc_sw128(w2_addr + (uint32_t)s2 * FFN_W2_STAGE_BYTES);
          const uint64_t db1 = umma_smem_desc_sw128(w2_addr + (uint32_t)s2 * FFN_W2_STAGE_BYTES + FFN_W2_HALF_BYTES);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < GEMM_BLOCK_K / 16; ++k) {
              const uint32_t acc = (uint32_t)((jj | kb2 | k) != 0);
              umma_bf16_2sm(tmem_base + TMEM_Y, da + (uint64_t)(k * 2), db0 + (uint64_t)(k * 2), IDESC2, acc);
              umma_bf16_2sm(tmem_base + TMEM_Y + 192, da + (uint64_t)(k * 2), db1 + (uint64_t)(k * 2), IDESC2, acc);
            }
            umma_commit_2sm(&w2_empty[s2]);
          }
          __syncwarp();
          if (++s2 == FFN_W2_STAGES) { s2 = 0; ph2 ^= 1; }
        }
        if (elect_one()) umma_commit_2sm(hs_free);         // Hs may be overwritten once these MMAs are done
        __syncwarp();
        ++g2c;
      };
      for (int mb = group; mb < m_blocks; mb += n_groups, ++mbi) {
        for (int j = 0; j < n_slices; ++j) {
          mbar_wait(h_free, (g1c & 1) ^ 1);                // the epilogue warps have drained the previous slice's H
          tc_fence_after();
          for (int kb = 0; kb < FFN_KB1; ++kb) {
            if (j == 0) mbar_wait(&x_full[kb], mbi & 1);
            mbar_wait(&w1_full[s1], ph1);
            tc_fence_after();
            const uint64_t da = umma_smem_desc_sw128(x_addr + (uint32_t)kb * GEMM_A_STAGE_BYTES);
            const uint64_t db = umma_smem_desc_sw128(w1_addr + (uint32_t)s1 * FFN_W1_STAGE_BYTES);
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < GEMM_BLOCK_K / 16; ++k)
                umma_bf16_2sm(tmem_base + TMEM_H, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), IDESC1, (uint32_t)((kb | k) != 0));
              umma_commit_2sm(&w1_empty[s1]);
              if (j == n_slices - 1) umma_commit_2sm(&x_empty[kb]);   // last use of this k-block of the slab
            }
            __syncwarp();
            if (++s1 == FFN_W1_STAGES) { s1 = 0; ph1 ^= 1; }
          }
          if (elect_one()) umma_commit_2sm(h_full);
          __syncwarp();
          ++g1c;
          if (j > 0) issue_g2(j - 1);
        }
        issue_g2(n_slices - 1);
        if (elect_one()) umma_commit_2sm(y_full);
        __syncwarp();
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue warps: TMEM lanes [32 wq, +32); column half c2 of H, column chunks c2, c2+2, c2+4 of Y =====
    const int e = warp - 4;
    const int wq = e & 3, c2 = e >> 2;
    const int row = wq * 32 + lane;
    const int sw = lane & 7;                    // 128-byte swizzle: 16-byte chunk ch of row r lives at ch ^ (r % 8)
    const uint32_t tlane = tmem_base + ((uint32_t)(wq * 32) << 16);
    // this warp's 32 rows of k-block c2 of the Hs tile == its 4 KB staging box in the final epilogue
    uint8_t* box = hs + (size_t)e * GEMM_STAGING_BYTES;
    const uint32_t s_row = smem_u32(box) + lane * 128;
    const uint32_t s_vecs = smem_u32(vecs), s_xs = smem_u32(xs);
    uint64_t* rbar = &resid_bar[e];
    const float inv_n = 1.0f / (float)FFN_H;
    uint32_t e1c = 0, mbi = 0, rcount = 0;
    for (int mb = group; mb < m_blocks; mb += n_groups, ++mbi) {
      // ---- E1: one slice of the intermediate at a time: H -> + b1 -> GELU -> bf16 -> Hs ----
#pragma unroll 1
      for (int j = 0; j < n_slices; ++j, ++e1c) {
        mbar_wait(h_full, e1c & 1);
        tc_fence_after();
        uint32_t v[64];
        tmem_ld64(tlane + TMEM_H + c2 * 64, v);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_leader(h_free);           // G1 of the next slice may overwrite H
        const float4* b1v = reinterpret_cast<const float4*>(p.b1 + j * FFN_SLICE + c2 * 64);
        uint4 o[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float4 ba = __ldg(b1v + 2 * q), bb = __ldg(b1v + 2 * q + 1);
          const float f0 = gelu_erf(__uint_as_float(v[q * 8 + 0]) + ba.x), f1 = gelu_erf(__uint_as_float(v[q * 8 + 1]) + ba.y);
          const float f2 = gelu_erf(__uint_as_float(v[q * 8 + 2]) + ba.z), f3 = gelu_erf(__uint_as_float(v[q * 8 + 3]) + ba.w);
          const float f4 = gelu_erf(__uint_as_float(v[q * 8 + 4]) + bb.x), f5 = gelu_erf(__uint_as_float(v[q * 8 + 5]) + bb.y);
          const float f6 = gelu_erf(__uint_as_float(v[q * 8 + 6]) + bb.z), f7 = gelu_erf(__uint_as_float(v[q * 8 + 7]) + bb.w);
          __nv_bfloat162 t0 = __floats2bfloat162_rn(f0, f1), t1 = __floats2bfloat162_rn(f2, f3);
          __nv_bfloat162 t2 = __floats2bfloat162_rn(f4, f5), t3 = __floats2bfloat162_rn(f6, f7);
          o[q].x = *reinterpret_cast<uint32_t*>(&t0); o[q].y = *reinterpret_cast<uint32_t*>(&t1);
          o[q].z = *reinterpret_cast<uint32_t*>(&t2); o[q].w = *reinterpret_cast<uint32_t*>(&t3);
        }
        mbar_wait(hs_free, (e1c & 1) ^ 1);                   // G2 of the previous slice has read Hs
#pragma unroll
        for (int q = 0; q < 8; ++q) sts128(s_row + ((q ^ sw) << 4), o[q]);
        fence_proxy_async();                                 // generic-proxy writes -> visible to the tensor core
        __syncwarp();
        if (lane == 0) mbar_arrive_leader(hs_full);
      }
      // ---- E2: y = Y + b2 + x (residual), LayerNorm, bf16, store over x ----
      const int m0 = (p.reverse ? m_blocks - 1 - mb : mb) * (GEMM_BLOCK_M * CG) + rank * GEMM_BLOCK_M + wq * 32;
      mbar_wait(y_full, mbi & 1);                            // every MMA of the row block is done (Hs is free too)
      tc_fence_after();
      float sum = 0.f, sumsq = 0.f;
#pragma unroll 1
      for (int t = 0; t < 3; ++t, ++rcount) {
        const int n0 = (c2 + 2 * t) * GEMM_OUT_CHUNK;
        __syncwarp();                                        // every lane is done reading the box
        if (lane == 0) {
          tma_store_wait_read();
          mbar_arrive_expect_tx(rbar, GEMM_STAGING_BYTES);
          tma_load_2d(box, &map_xio, rbar, n0, m0);
        }
        __syncwarp();
        uint32_t v[64];
        tmem_ld64(tlane + TMEM_Y + n0, v);
        tmem_ld_wait();
        mbar_wait(rbar, rcount & 1);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const uint4 b0 = lds128(s_vecs + (n0 + q * 8) * 4), b1 = lds128(s_vecs + (n0 + q * 8 + 4) * 4);
          const uint4 r = lds128(s_row + ((q ^ sw) << 4));
          const __nv_bfloat162* r2 = reinterpret_cast<const __nv_bfloat162*>(&r);
          const float bb[8] = {__uint_as_float(b0.x), __uint_as_float(b0.y), __uint_as_float(b0.z), __uint_as_float(b0.w),
                               __uint_as_float(b1.x), __uint_as_float(b1.y), __uint_as_float(b1.z), __uint_as_float(b1.w)};
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float y0 = __uint_as_float(v[q * 8 + 2 * i]) + bb[2 * i] + __low2float(r2[i]);
            const float y1 = __uint_as_float(v[q * 8 + 2 * i + 1]) + bb[2 * i + 1] + __high2float(r2[i]);
            sum += y0 + y1;
            sumsq = fmaf(y0, y0, sumsq);
            sumsq = fmaf(y1, y1, sumsq);
            v[q * 8 + 2 * i] = __float_as_uint(y0);
            v[q * 8 + 2 * i + 1] = __float_as_uint(y1);
          }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint32_t w[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) w[i] = v[q * 16 + i];
          tmem_st16(tlane + TMEM_Y + n0 + q * 16, w);
        }
      }
      tmem_st_wait();
      const uint32_t xs_par = s_xs + (mbi & 1) * (2 * 128 * 8);
      sts32(xs_par + (c2 * 128 + row) * 8, __float_as_uint(sum));
      sts32(xs_par + (c2 * 128 + row) * 8 + 4, __float_as_uint(sumsq));
      named_bar_sync(1 + wq, 64);                            // the two warps of this lane group
      const float tot = __uint_as_float(lds32(xs_par + row * 8)) + __uint_as_float(lds32(xs_par + (128 + row) * 8));
      const float totsq = __uint_as_float(lds32(xs_par + row * 8 + 4)) + __uint_as_float(lds32(xs_par + (128 + row) * 8 + 4));
      const float mean = tot * inv_n;
      const float rstd = rsqrtf(fmaxf(totsq * inv_n - mean * mean, 0.f) + p.eps);
      const float shift = -mean * rstd;
#pragma unroll 1
      for (int t = 0; t < 3; ++t) {
        const int n0 = (c2 + 2 * t) * GEMM_OUT_CHUNK;
        uint32_t v[64];
        tmem_ld64(tlane + TMEM_Y + n0, v);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (t == 2) mbar_arrive_leader(y_free);            // this warp's share of Y is in registers
          tma_store_wait_read();                             // the previous store has finished reading the box
        }
        __syncwarp();
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const uint4 g0 = lds128(s_vecs + (FFN_H + n0 + q * 8) * 4), g1 = lds128(s_vecs + (FFN_H + n0 + q * 8 + 4) * 4);
          const uint4 e0 = lds128(s_vecs + (2 * FFN_H + n0 + q * 8) * 4), e1 = lds128(s_vecs + (2 * FFN_H + n0 + q * 8 + 4) * 4);
          const float gg[8] = {__uint_as_float(g0.x), __uint_as_float(g0.y), __uint_as_float(g0.z), __uint_as_float(g0.w),
                               __uint_as_float(g1.x), __uint_as_float(g1.y), __uint_as_float(g1.z), __uint_as_float(g1.w)};
          const float ee[8] = {__uint_as_float(e0.x), __uint_as_float(e0.y), __uint_as_float(e0.z), __uint_as_float(e0.w),
                               __uint_as_float(e1.x), __uint_as_float(e1.y), __uint_as_float(e1.z), __uint_as_float(e1.w)};
          float f[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] = fmaf(fmaf(__uint_as_float(v[q * 8 + i]), rstd, shift), gg[i], ee[i]);
          uint4 o;
          __nv_bfloat162 t0 = __floats2bfloat162_rn(f[0], f[1]), t1 = __floats2bfloat162_rn(f[2], f[3]);
          __nv_bfloat162 t2 = __floats2bfloat162_rn(f[4], f[5]), t3 = __floats2bfloat162_rn(f[6], f[7]);
          o.x = *reinterpret_cast<uint32_t*>(&t0); o.y = *reinterpret_cast<uint32_t*>(&t1);
          o.z = *reinterpret_cast<uint32_t*>(&t2); o.w = *reinterpret_cast<uint32_t*>(&t3);
          sts128(s_row + ((q ^ sw) << 4), o);
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&map_xio, box, n0, m0);
          tma_store_commit();
        }
      }
      // the box is this warp's part of Hs: the next row block's first slice may only overwrite it once the
      // last store has read it
      if (lane == 0) tma_store_wait_read();
      __syncwarp();
    }
    if (lane == 0) tma_store_wait_all();
  }

  tc_fence_before();
  cluster_sync_all();                          // no CTA leaves while its peer may still signal its barriers
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_2sm(tmem_base, 512);
  }
}

}  // namespace b200
