// encoder_kernels.cuh -- device code for the chunk-embedding path.
//
// Restates what SentenceTransformer.encode does for a BERT-class model
// (sentence_transformer_provider.py:93; pipeline in oracle/encode_oracle.py):
//
//   embed_ln_kernel        K1   word + position + type embeddings -> LayerNorm
//   gemm_bf16_kernel       K2/K4/K5/K6  X[M,K] . W[N,K]^T on tcgen05 (TMA-fed, TMEM accumulators,
//                               persistent, warp-specialised) with fused epilogues:
//                               +bias | +bias,GELU(erf) | +bias,+residual
//   layernorm_kernel       (K4/K6 tail) LayerNorm over the hidden dimension
//   attention_kernel       K3   softmax(QK^T/sqrt(dh) + key-padding mask) V, flash-style
//   pool_normalize_kernel  K7   masked mean over tokens + L2 normalise
//
// Activations are bf16 [tokens, features] row-major; accumulation, LayerNorm statistics, softmax
// and pooling are fp32.
#pragma once
#include "common.cuh"

namespace b200 {

// =============================================================================================
// K1: embeddings + LayerNorm.  One warp per token; H = 128 * NV elements, 4 per lane per pass.
// =============================================================================================
B200_DEVINL void bf16x4_to_f32(uint2 u, float (&f)[4]) {
  const __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&u.x);
  const __nv_bfloat162 b = *reinterpret_cast<__nv_bfloat162*>(&u.y);
  f[0] = __low2float(a); f[1] = __high2float(a); f[2] = __low2float(b); f[3] = __high2float(b);
}
B200_DEVINL uint2 f32x4_to_bf16(const float (&f)[4]) {
  __nv_bfloat162 a = __floats2bfloat162_rn(f[0], f[1]);
  __nv_bfloat162 b = __floats2bfloat162_rn(f[2], f[3]);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&a);
  u.y = *reinterpret_cast<uint32_t*>(&b);
  return u;
}

// LayerNorm of one row held as NV x 4 floats per lane (two-pass, fp32), affine, bf16 out
template <int NV>
B200_DEVINL void warp_layernorm_store(float (&v)[NV][4], int H, float eps, const float* __restrict__ gamma,
                                      const float* __restrict__ beta, __nv_bfloat16* __restrict__ out, int lane) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) s += (v[i][0] + v[i][1]) + (v[i][2] + v[i][3]);
  const float mean = warp_sum(s) / (float)H;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float d = v[i][j] - mean;
      q = fmaf(d, d, q);
    }
  const float rstd = rsqrtf(warp_sum(q) / (float)H + eps);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = i * 128 + lane * 4;
    const float4 g = *reinterpret_cast<const float4*>(gamma + c);
    const float4 b = *reinterpret_cast<const float4*>(beta + c);
    float o[4];
    o[0] = (v[i][0] - mean) * rstd * g.x + b.x;
    o[1] = (v[i][1] - mean) * rstd * g.y + b.y;
    o[2] = (v[i][2] - mean) * rstd * g.z + b.z;
    o[3] = (v[i][3] - mean) * rstd * g.w + b.w;
    *reinterpret_cast<uint2*>(out + c) = f32x4_to_bf16(o);
  }
}

template <int NV>
__global__ void __launch_bounds__(256) embed_ln_kernel(const int32_t* __restrict__ ids, int n_tokens, int S, int vocab,
                                                       const __nv_bfloat16* __restrict__ word,
                                                       const __nv_bfloat16* __restrict__ pos,
                                                       const __nv_bfloat16* __restrict__ type0,
                                                       const float* __restrict__ gamma, const float* __restrict__ beta,
                                                       float eps, __nv_bfloat16* __restrict__ x) {
  constexpr int H = NV * 128;
  const int tok = blockIdx.x * (blockDim.x / kWarp) + (threadIdx.x / kWarp);
  if (tok >= n_tokens) return;
  const int lane = threadIdx.x & 31;
  int id = ids[tok];
  id = min(max(id, 0), vocab - 1);
  const int t = tok % S;
  float v[NV][4];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = i * 128 + lane * 4;
    float w[4], p[4], ty[4];
    bf16x4_to_f32(__ldg(reinterpret_cast<const uint2*>(word + (size_t)id * H + c)), w);
    bf16x4_to_f32(__ldg(reinterpret_cast<const uint2*>(pos + (size_t)t * H + c)), p);
    bf16x4_to_f32(__ldg(reinterpret_cast<const uint2*>(type0 + c)), ty);
#pragma unroll
    for (int j = 0; j < 4; ++j) v[i][j] = (w[j] + p[j]) + ty[j];
  }
  warp_layernorm_store<NV>(v, H, eps, gamma, beta, x + (size_t)tok * H, lane);
}

// Deferred LayerNorm (see EPI_ROWFOLD): statistic slots that make a row read as "already normalised" -- sum 0, sum of
// squares = width, so mean = 0 and rstd = 1 (to 1e-12).  The embedding LayerNorm is NOT deferred: the raw embedding sum
// of a model with a large common offset would lose its signal in bf16, so layer 0 reads embed_ln_kernel's output through
// these slots with unfolded weights and gamma = 1, beta = 0.
__global__ void ln_identity_stats_kernel(float* __restrict__ stats, int64_t n_rows, int parts, float width) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_rows * parts) return;
  reinterpret_cast<float2*>(stats)[i] = (i % parts) == 0 ? make_float2(0.f, width) : make_float2(0.f, 0.f);
}

template <int NV>
__global__ void __launch_bounds__(256) layernorm_kernel(const __nv_bfloat16* __restrict__ y, int n_rows,
                                                        const float* __restrict__ gamma, const float* __restrict__ beta,
                                                        float eps, __nv_bfloat16* __restrict__ x) {
  constexpr int H = NV * 128;
  const int row = blockIdx.x * (blockDim.x / kWarp) + (threadIdx.x / kWarp);
  if (row >= n_rows) return;
  const int lane = threadIdx.x & 31;
  float v[NV][4];
#pragma unroll
  for (int i = 0; i < NV; ++i)
    bf16x4_to_f32(*reinterpret_cast<const uint2*>(y + (size_t)row * H + i * 128 + lane * 4), v[i]);
  warp_layernorm_store<NV>(v, H, eps, gamma, beta, x + (size_t)row * H, lane);
}

// =============================================================================================
// GEMM: out[M,N] = epilogue( A[M,K] . W[N,K]^T ), bf16 in, fp32 accumulate in TMEM, bf16 out.
// Persistent CTAs (one per SM) walk 128 x BLOCK_N output tiles; roles:
//   warp 0 TMA producer | warp 1 MMA issuer | warp 2 TMEM allocator | warps 4.. epilogue
// Two TMEM accumulator stages let tile i's epilogue overlap tile i+1's MMAs.
//
// Epilogue: one warp per (32-row TMEM lane group, 64-column chunk).  The warp pulls its 32x64 fp32
// block with ONE tcgen05.ld (x64), hands the accumulator stage back to the MMA warp, applies
// bias / GELU / residual, packs to bf16 into a 4 KB shared-memory box in the 128-byte-swizzled
// layout and stores it with ONE TMA store: fully coalesced 128-byte rows, no per-thread global
// address generation.  The residual tile arrives the same way (TMA load into the same box, added in
// place).  (A first version stored 16 bytes per thread straight to global memory: 32 different
// cache lines per instruction, which made the L1/LSU wavefront rate -- not the MMA -- the bound.)
// =============================================================================================
constexpr int GEMM_BLOCK_M = 128;
constexpr int GEMM_BLOCK_K = 64;
constexpr int GEMM_A_STAGE_BYTES = GEMM_BLOCK_M * GEMM_BLOCK_K * 2;
constexpr int GEMM_OUT_CHUNK = 64;                                  // columns per epilogue warp = 128 B of bf16
constexpr int GEMM_STAGING_BYTES = 32 * GEMM_OUT_CHUNK * 2;         // 32 rows x 128 B per epilogue warp

constexpr int gemm_epi_warps(int block_n) { return 4 * (block_n / GEMM_OUT_CHUNK); }
constexpr int gemm_threads(int block_n) { return 128 + 32 * gemm_epi_warps(block_n); }

// Epilogue of the GEMM kernels, a bit set: every epilogue adds the bias.
//   EPI_BIAS_GELU   GELU on the result (FFN1)
//   EPI_BIAS_RESID  + the residual tile (out-projection, FFN2)
//   EPI_HEADNORM    the output is the QKV projection of 64-wide heads: besides storing it, fold max_r |q_r|^2 and
//                   max_j |k_j|^2 of every (sequence, head) into GemmParams::hn_q2 / hn_k2 (what att_norms_kernel
//                   would re-read the whole Q and K for).  An epilogue warp holds 32 rows x 64 columns = one head's slice of
//                   32 tokens, one row per thread, so the squared norm is a thread-local sum and the maximum one warp
//                   reduction + one atomicMax (non-negative floats order like their bit patterns).
//
// Deferred LayerNorm (hidden sizes without the fused GEMM + LayerNorm kernel, e.g. 768): x = LN(y) is never
// written.  A producer stores y (bf16) and, per row and 64-column chunk, (sum, sum of squares) of what it stored
// (EPI_STATS: plain stores into the row's slot for that chunk -- no atomics, deterministic).  Consumers rebuild
// what they need from y and those sums:
//   EPI_ROWFOLD   the A operand is y, the weights are W' = W (.) gamma (column k scaled by gamma_k), and
//                   x W^T = rstd (y W'^T - mean c) + W beta,   c_n = sum_k W'_nk
//                 so the epilogue computes rstd_r (acc - mean_r c_n) + b'_n with b' = b + W beta folded on the host.
//   EPI_RESID_LN  the residual tile holds y: x = (y - mean) rstd gamma + beta is formed on the fly (implies RESID).
enum { EPI_BIAS = 0, EPI_BIAS_GELU = 1, EPI_BIAS_RESID = 2, EPI_HEADNORM = 4, EPI_ROWFOLD = 8, EPI_RESID_LN = 16, EPI_STATS = 32 };
constexpr int gemm_epi_consts(int epi) { return (epi & (EPI_ROWFOLD | EPI_RESID_LN)) ? 3 : 1; }   // 64-float slices staged per warp

struct GemmParams {
  int M, N, K;         // M multiple of 128, N multiple of BLOCK_N, K multiple of 64
  int nstage;
  const float* bias;   // [N]
  const float* gamma;  // [N]  LayerNorm weight (gemm_ln_kernel only)
  const float* beta;   // [N]  LayerNorm bias   (gemm_ln_kernel only)
  float eps;
  int reverse;         // 1: walk the row blocks from the last to the first.  Consecutive kernels of a layer
                       // alternate, so each starts with the rows its producer wrote LAST -- the part of a
                       // 300-400 MB activation that is still in the 126 MB L2.
  // EPI_HEADNORM: [batch * heads] running maxima (zeroed before the launch), sequence length, hidden size, heads
  float* hn_q2;
  float* hn_k2;
  int hn_S, hn_H, hn_heads, hn_rows;   // hn_rows = batch * S: rows that belong to a sequence
  // deferred LayerNorm: statistic slots [M][ln_parts][2] of the tensor being normalised (ROWFOLD: the A operand,
  // RESID_LN: the residual), 1 / its width, the folded column sums c[N] (ROWFOLD), gamma / beta above (RESID_LN);
  // EPI_STATS: the slots of the tensor being produced ([M][N / 64][2])
  const float* ln_stats_in;
  int ln_parts;
  float ln_inv_n;
  const float* fold_c;
  float* ln_stats_out;
};

// GELU(x) = x * Phi(x) (the exact-erf form BERT uses).  Phi is evaluated as
//   0.5 * (1 + tanh(x * (a0 + a1 x^2 + a2 x^4)))
// with a0..a2 fitted (minimax over |x| <= 6) to the erf form: max |error| of the result is 1.3e-4,
// a thirtieth of the bf16 rounding step of the stored activation.  One MUFU (tanh.approx) and six
// FMA-pipe instructions per element: the erf series with exp + reciprocal costs two MUFU ops, and at
// 4 SFU lanes per SM sub-partition that made this epilogue, not the MMA, the FFN1 bottleneck.
B200_DEVINL float gelu_erf(float x) {
  const float x2 = fminf(x * x, 36.0f);   // beyond |x| = 6 the tanh argument is saturated either way
  const float p = fmaf(x2, fmaf(x2, -3.9992863e-4f, 3.7419684e-2f), 7.9673827e-1f);
  float th;
  asm("tanh.approx.f32 %0, %1;" : "=f"(th) : "f"(x * p));
  const float hx = 0.5f * x;
  return fmaf(hx, th, hx);
}

#ifdef B200_GEMM_TIMING
// Development aid (not compiled into the shipped library): where the MMA issuer of CTA pair 0 spends its
// clocks.  [0] waiting for a free accumulator, [1] waiting for operands, [2] issuing, [3] tiles, [4] total.
__device__ unsigned long long g_gemm_timing[8];
#endif

// One epilogue warp's share of one output tile: 32 rows (TMEM lanes) x 64 columns.  Accumulator ->
// registers (one tcgen05.ld.x64), bias (+ GELU | + residual, which arrives by TMA in the staging box),
// bf16, 128-byte-swizzled staging box, TMA store.
template <int EPI, int CG>
B200_DEVINL void gemm_epi_tile(const CUtensorMap& map_out, const CUtensorMap& map_resid, const GemmParams& p, int m0, int n0,
                               uint8_t* stg, float* sbias, uint64_t* rbar, uint8_t* my_row, int sw, int lane, uint32_t tacc,
                               uint64_t* acc_full_bar, uint32_t aph, uint64_t* acc_empty_bar, uint32_t it) {
  const float* bias = p.bias;
  float head_sq = 0.f;       // EPI_HEADNORM: |row slice|^2 of the values as stored (bf16)
  __syncwarp();
  sts32(smem_u32(sbias) + lane * 4, __float_as_uint(__ldg(bias + n0 + lane)));
  sts32(smem_u32(sbias) + (lane + 32) * 4, __float_as_uint(__ldg(bias + n0 + 32 + lane)));
  float r_mean = 0.f, r_rstd = 1.f;   // deferred LayerNorm: this thread's row
  if (EPI & (EPI_ROWFOLD | EPI_RESID_LN)) {
    const float* c1 = (EPI & EPI_ROWFOLD) ? p.fold_c : p.gamma;
    sts32(smem_u32(sbias) + (64 + lane) * 4, __float_as_uint(__ldg(c1 + n0 + lane)));
    sts32(smem_u32(sbias) + (96 + lane) * 4, __float_as_uint(__ldg(c1 + n0 + 32 + lane)));
    if (EPI & EPI_RESID_LN) {
      sts32(smem_u32(sbias) + (128 + lane) * 4, __float_as_uint(__ldg(p.beta + n0 + lane)));
      sts32(smem_u32(sbias) + (160 + lane) * 4, __float_as_uint(__ldg(p.beta + n0 + 32 + lane)));
    }
    const float4* sp = reinterpret_cast<const float4*>(p.ln_stats_in + (size_t)(m0 + lane) * 2 * p.ln_parts);
    float s1 = 0.f, s2 = 0.f;
    for (int i = 0; i < p.ln_parts / 2; ++i) {
      const float4 t = __ldg(sp + i);
      s1 += t.x + t.z;
      s2 += t.y + t.w;
    }
    r_mean = s1 * p.ln_inv_n;
    r_rstd = rsqrtf(fmaxf(fmaf(-r_mean, r_mean, s2 * p.ln_inv_n), 0.f) + p.eps);
  }
  float st1 = 0.f, st2 = 0.f;         // EPI_STATS: sums of this thread's 64 stored values
  if (lane == 0) {
    tma_store_wait_read();   // the previous tile's store has finished reading this warp's box
    if (EPI & EPI_BIAS_RESID) {
      mbar_arrive_expect_tx(rbar, GEMM_STAGING_BYTES);
      tma_load_2d(stg, &map_resid, rbar, n0, m0);
    }
  }
  __syncwarp();
  mbar_wait(acc_full_bar, aph);
  tc_fence_after();
  uint32_t v[64];
  tmem_ld64(tacc, v);
  tmem_ld_wait();
  tc_fence_before();         // this warp's share of the accumulator stage is in registers
  __syncwarp();
  if (lane == 0) {
    if (CG == 2) mbar_arrive_leader(acc_empty_bar); else mbar_arrive(acc_empty_bar);
  }
  if (EPI & EPI_BIAS_RESID) mbar_wait(rbar, it & 1);
#ifdef B200_GEMM_TIMING
  const unsigned long long dbg = g_gemm_timing[7];
  if (dbg == 1) return;   // experiment: accumulators drained, nothing computed or stored
#endif
  const uint32_t s_bias = smem_u32(sbias), s_row = smem_u32(my_row);   // shared-space accesses (see lds128)
#pragma unroll
  for (int j = 0; j < 8; ++j) {   // 8 columns = one 16-byte piece of this thread's row
    const uint4 b0 = lds128(s_bias + j * 32), b1 = lds128(s_bias + j * 32 + 16);
    const uint32_t bw[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
    float f[8];
    if (EPI & EPI_ROWFOLD) {
      const uint4 c0 = lds128(s_bias + 256 + j * 32), c1 = lds128(s_bias + 256 + j * 32 + 16);
      const uint32_t cw[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
        f[i] = fmaf(r_rstd, fmaf(-r_mean, __uint_as_float(cw[i]), __uint_as_float(v[j * 8 + i])), __uint_as_float(bw[i]));
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(v[j * 8 + i]) + __uint_as_float(bw[i]);
    }
    if (EPI & EPI_BIAS_GELU) {
#pragma unroll
      for (int i = 0; i < 8; ++i) f[i] = gelu_erf(f[i]);
    }
    const uint32_t slot = s_row + ((j ^ sw) << 4);
    if (EPI & EPI_BIAS_RESID) {
      const uint4 r = lds128(slot);
      const __nv_bfloat162* r2 = reinterpret_cast<const __nv_bfloat162*>(&r);
      if (EPI & EPI_RESID_LN) {
        const uint4 g0 = lds128(s_bias + 256 + j * 32), g1 = lds128(s_bias + 256 + j * 32 + 16);
        const uint4 e0 = lds128(s_bias + 512 + j * 32), e1 = lds128(s_bias + 512 + j * 32 + 16);
        const uint32_t gw[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
        const uint32_t ew[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float a0 = r_rstd * __uint_as_float(gw[2 * i]), a1 = r_rstd * __uint_as_float(gw[2 * i + 1]);
          f[2 * i] += fmaf(__low2float(r2[i]) - r_mean, a0, __uint_as_float(ew[2 * i]));
          f[2 * i + 1] += fmaf(__high2float(r2[i]) - r_mean, a1, __uint_as_float(ew[2 * i + 1]));
        }
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          f[2 * i] += __low2float(r2[i]);
          f[2 * i + 1] += __high2float(r2[i]);
        }
      }
    }
    uint4 o;
    __nv_bfloat162 t0 = __floats2bfloat162_rn(f[0], f[1]);
    __nv_bfloat162 t1 = __floats2bfloat162_rn(f[2], f[3]);
    __nv_bfloat162 t2 = __floats2bfloat162_rn(f[4], f[5]);
    __nv_bfloat162 t3 = __floats2bfloat162_rn(f[6], f[7]);
    o.x = *reinterpret_cast<uint32_t*>(&t0); o.y = *reinterpret_cast<uint32_t*>(&t1);
    o.z = *reinterpret_cast<uint32_t*>(&t2); o.w = *reinterpret_cast<uint32_t*>(&t3);
    sts128(slot, o);
    if (EPI & (EPI_HEADNORM | EPI_STATS)) {
      const __nv_bfloat162 tt[4] = {t0, t1, t2, t3};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float lo = __low2float(tt[i]), hi = __high2float(tt[i]);
        if (EPI & EPI_HEADNORM) head_sq = fmaf(lo, lo, fmaf(hi, hi, head_sq));
        if (EPI & EPI_STATS) {
          st1 += lo + hi;
          st2 = fmaf(lo, lo, fmaf(hi, hi, st2));
        }
      }
    }
  }
  if (EPI & EPI_STATS)
    *reinterpret_cast<float2*>(p.ln_stats_out + ((size_t)(m0 + lane) * (p.N / GEMM_OUT_CHUNK) + n0 / GEMM_OUT_CHUNK) * 2) =
        make_float2(st1, st2);
  if (EPI & EPI_HEADNORM) {
    if (n0 < 2 * p.hn_H && m0 < p.hn_rows) {   // Q and K columns (V needs no norm) of real tokens (rows past the batch pad the
                                               // last tile and belong to no sequence); warp-uniform
      const float mx = warp_max(head_sq);
      if (lane == 0) {
        const int is_k = n0 >= p.hn_H;
        const int h = (n0 - is_k * p.hn_H) / GEMM_OUT_CHUNK;
        atomicMax(reinterpret_cast<unsigned int*>(is_k ? p.hn_k2 : p.hn_q2) + (m0 / p.hn_S) * p.hn_heads + h, __float_as_uint(mx));
      }
    }
  }
#ifdef B200_GEMM_TIMING
  if (dbg == 3) return;      // experiment: no fence, no store
#endif
  fence_proxy_async();       // generic-proxy writes -> visible to the TMA (async proxy)
  __syncwarp();
#ifdef B200_GEMM_TIMING
  if (dbg == 2) return;      // experiment: no store
#endif
  if (lane == 0) {
#ifdef B200_GEMM_TIMING
    if (dbg == 4) m0 &= 255;   // experiment: every store lands in the first 256 rows (stays in L2)
#endif
    tma_store_2d(&map_out, stg, n0, m0);
    tma_store_commit();
  }
}

// CG = 1: one CTA per 128 x BLOCK_N tile.  CG = 2: a CTA pair (cluster of 2, cta_group::2) per
// 256 x BLOCK_N tile -- each CTA loads its own 128 rows of A but only HALF of the weight tile, and the
// leader's single MMA drives both SMs' tensor cores.  That halves the weight bytes every SM pulls
// through L2 per flop (the CG = 1 kernel saturates the L2 -> SM path at ~77 flop/B) and leaves room
// for a deeper TMA ring.
template <int BLOCK_N, int EPI, int CG>
__global__ void __launch_bounds__(gemm_threads(BLOCK_N), 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 const __grid_constant__ CUtensorMap map_out, const __grid_constant__ CUtensorMap map_resid,
                 const GemmParams p) {
  constexpr uint32_t B_ROWS = BLOCK_N / CG;                      // weight rows this CTA loads per stage
  constexpr uint32_t B_STAGE_BYTES = B_ROWS * GEMM_BLOCK_K * 2;
  constexpr uint32_t STAGE_BYTES = GEMM_A_STAGE_BYTES + B_STAGE_BYTES;
  constexpr uint32_t IDESC = umma_idesc_bf16_f32(GEMM_BLOCK_M * CG, BLOCK_N);
  constexpr uint32_t TMEM_COLS = (2 * BLOCK_N <= 128) ? 128 : ((2 * BLOCK_N <= 256) ? 256 : 512);
  constexpr int NEPI = gemm_epi_warps(BLOCK_N);
  static_assert(BLOCK_N % 64 == 0 && BLOCK_N <= 256, "BLOCK_N must be 64..256 in steps of 64");
  static_assert(CG == 1 || CG == 2, "CTA group is 1 or 2");
  static_assert(STAGE_BYTES % 1024 == 0, "stages must keep the 1024-byte swizzle alignment");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* staging = smem + (size_t)p.nstage * STAGE_BYTES;                 // [NEPI][4 KB], 1024-aligned
  uint64_t* bars = reinterpret_cast<uint64_t*>(staging + (size_t)NEPI * GEMM_STAGING_BYTES);
  uint64_t* full = bars;                       // [nstage]  TMA -> MMA          (leader's copy is the one used)
  uint64_t* empty = bars + p.nstage;           // [nstage]  MMA -> TMA          (one per CTA)
  uint64_t* acc_full = bars + 2 * p.nstage;    // [2]       MMA -> epilogue     (one per CTA)
  uint64_t* acc_empty = acc_full + 2;          // [2]       epilogue -> MMA     (leader's copy is the one used)
  uint64_t* resid_bar = acc_empty + 2;         // [NEPI]    residual box landed (one per epilogue warp)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(resid_bar + NEPI);
  float* bias_smem = reinterpret_cast<float*>(tmem_slot + 4);               // [NEPI][gemm_epi_consts(EPI)][64]

  const int warp = uniform_warp_idx();
  const int lane = threadIdx.x & 31;
  const int rank = (CG == 2) ? (int)cluster_ctarank() : 0;
  const bool leader = rank == 0;
  const int n_tiles = p.N / BLOCK_N;
  const int total_tiles = (p.M / (GEMM_BLOCK_M * CG)) * n_tiles;   // tiles of the CTA group
  const int group = blockIdx.x / CG, n_groups = gridDim.x / CG;
  const int num_k = p.K / GEMM_BLOCK_K;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    tma_prefetch_desc(&map_out);
    if (EPI & EPI_BIAS_RESID) tma_prefetch_desc(&map_resid);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < p.nstage; ++i) {
      mbar_init(&full[i], CG);                 // one arrival per CTA of the group
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], NEPI * CG);
    }
    for (int i = 0; i < NEPI; ++i) mbar_init(&resid_bar[i], 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    if (CG == 2) tmem_alloc_2sm(tmem_slot, TMEM_COLS);
    else tmem_alloc(tmem_slot, TMEM_COLS);
  }
  tc_fence_before();
  if (CG == 2) cluster_sync_all();             // the peer's barriers are initialised before anyone signals them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                                  // the producer of A / the residual has finished
  pdl_launch_dependents();

  if (warp == 0) {
    // ===== TMA producer (every CTA loads its own A rows and its share of the weight tile); the warp runs
    // the loop converged, one elected lane issues =====
    {
      int s = 0;
      uint32_t ph = 0;
      for (int tile = group; tile < total_tiles; tile += n_groups) {
        const int mt = tile / n_tiles, mtr = p.reverse ? total_tiles / n_tiles - 1 - mt : mt;
        const int m0 = mtr * (GEMM_BLOCK_M * CG) + rank * GEMM_BLOCK_M;
        const int n0 = (tile % n_tiles) * BLOCK_N + rank * (int)B_ROWS;
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(&empty[s], ph ^ 1);
          uint8_t* sa = smem + (size_t)s * STAGE_BYTES;
          if (elect_one()) {
            if (CG == 2) {
              if (leader) mbar_arrive_expect_tx(&full[s], STAGE_BYTES * CG);
              else mbar_arrive_leader(&full[s]);
              tma_load_2d_2sm(sa, &map_a, &full[s], kb * GEMM_BLOCK_K, m0);
              tma_load_2d_2sm(sa + GEMM_A_STAGE_BYTES, &map_b, &full[s], kb * GEMM_BLOCK_K, n0);
            } else {
              mbar_arrive_expect_tx(&full[s], STAGE_BYTES);
              tma_load_2d(sa, &map_a, &full[s], kb * GEMM_BLOCK_K, m0);
              tma_load_2d(sa + GEMM_A_STAGE_BYTES, &map_b, &full[s], kb * GEMM_BLOCK_K, n0);
            }
          }
          __syncwarp();
          if (++s == p.nstage) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the leader CTA's warp 1 runs the loop converged, one elected lane issues =====
    if (leader) {
      int s = 0;
      uint32_t ph = 0, it = 0;
      const uint32_t smem_addr = smem_u32(smem);
      for (int tile = group; tile < total_tiles; tile += n_groups, ++it) {
        const uint32_t as = it & 1, aph = (it >> 1) & 1;
#ifdef B200_GEMM_TIMING
        const long long g0 = clock64();
#endif
        mbar_wait(&acc_empty[as], aph ^ 1);
        tc_fence_after();
#ifdef B200_GEMM_TIMING
        long long g1 = clock64(), gw = 0;
        if (blockIdx.x == 0 && lane == 0) atomicAdd(&g_gemm_timing[0], (unsigned long long)(g1 - g0));
#endif
        const uint32_t tacc = tmem_base + as * BLOCK_N;
        for (int kb = 0; kb < num_k; ++kb) {
#ifdef B200_GEMM_TIMING
          const long long w0 = clock64();
#endif
          mbar_wait(&full[s], ph);
          tc_fence_after();
#ifdef B200_GEMM_TIMING
          gw += clock64() - w0;
#endif
          const uint32_t a_addr = smem_addr + (uint32_t)s * STAGE_BYTES;
          const uint64_t da = umma_smem_desc_sw128(a_addr);
          const uint64_t db = umma_smem_desc_sw128(a_addr + GEMM_A_STAGE_BYTES);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < GEMM_BLOCK_K / 16; ++k) {
              if (CG == 2) umma_bf16_2sm(tacc, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), IDESC, (uint32_t)((kb | k) != 0));
              else umma_bf16(tacc, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), IDESC, (uint32_t)((kb | k) != 0));
            }
            if (CG == 2) umma_commit_2sm(&empty[s]); else umma_commit(&empty[s]);
          }
          __syncwarp();
          if (++s == p.nstage) { s = 0; ph ^= 1; }
        }
        if (elect_one()) {
          if (CG == 2) umma_commit_2sm(&acc_full[as]); else umma_commit(&acc_full[as]);
        }
        __syncwarp();
#ifdef B200_GEMM_TIMING
        if (blockIdx.x == 0 && lane == 0) {
          const long long g2 = clock64();
          atomicAdd(&g_gemm_timing[1], (unsigned long long)gw);
          if (tile % n_tiles == 0) {      // first column tile of a row block: its A rows come from HBM
            atomicAdd(&g_gemm_timing[5], (unsigned long long)gw);
            atomicAdd(&g_gemm_timing[6], 1ull);
          }
          atomicAdd(&g_gemm_timing[2], (unsigned long long)(g2 - g1 - gw));
          atomicAdd(&g_gemm_timing[3], 1ull);
          atomicAdd(&g_gemm_timing[4], (unsigned long long)(g2 - g0));
        }
#endif
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue warp: TMEM lanes [32 wq, +32) x columns [64 chunk, +64) of this CTA's 128 rows =====
    const int e = warp - 4;
    const int wq = e & 3;         // == warp % 4: the TMEM lane group this warp may read
    const int chunk = e >> 2;
    uint8_t* stg = staging + (size_t)e * GEMM_STAGING_BYTES;
    float* sbias = bias_smem + e * GEMM_OUT_CHUNK * gemm_epi_consts(EPI);
    uint64_t* rbar = &resid_bar[e];
    uint8_t* my_row = stg + lane * 128;
    const int sw = lane & 7;       // 128-byte swizzle: 16-byte chunk j of row r lives at chunk j ^ (r % 8)
    uint32_t it = 0;
    for (int tile = group; tile < total_tiles; tile += n_groups, ++it) {
      const uint32_t as = it & 1, aph = (it >> 1) & 1;
      const int mt = tile / n_tiles, mtr = p.reverse ? total_tiles / n_tiles - 1 - mt : mt;
      const int m0 = mtr * (GEMM_BLOCK_M * CG) + rank * GEMM_BLOCK_M + wq * 32;
      const int n0 = (tile % n_tiles) * BLOCK_N + chunk * GEMM_OUT_CHUNK;
      gemm_epi_tile<EPI, CG>(map_out, map_resid, p, m0, n0, stg, sbias, rbar, my_row, sw, lane,
                             tmem_base + ((uint32_t)(wq * 32) << 16) + as * BLOCK_N + chunk * GEMM_OUT_CHUNK,
                             &acc_full[as], aph, &acc_empty[as], it);
    }
    if (lane == 0) tma_store_wait_all();
  }

  tc_fence_before();
  if (CG == 2) cluster_sync_all();             // no CTA leaves while its peer may still signal its barriers
  else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    if (CG == 2) tmem_dealloc_2sm(tmem_base, TMEM_COLS);
    else tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// A-stationary variant for K <= GEMM_AST_MAX_KB * 64 (the encoder's K = hidden GEMMs), CTA pairs only.
// The generic kernel re-reads the activation tile once per column tile (N / BLOCK_N times: 6 for the
// QKV projection, 8 for FFN1) and at ~10 TB/s of L2 -> SM traffic that, not the tensor pipe, is what
// holds it at 60 % MMA duty.  Here a CTA keeps its 128 x K slab of A in shared memory for ALL column
// tiles of its row block and streams only the weight half-tiles through a ring:
//   L2 -> SM bytes, QKV projection: 1.06 GB -> 0.55 GB;  FFN1: 1.41 GB -> 0.70 GB.
// Each 64-wide k-block of the slab has its own full / empty barrier, so the producer refills the
// slab with the next row block k-block by k-block while the last column tile is still being issued.
constexpr int GEMM_AST_MAX_KB = 6;

template <int BLOCK_N, int EPI>
__global__ void __launch_bounds__(gemm_threads(BLOCK_N), 1)
gemm_bf16_ast_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                     const __grid_constant__ CUtensorMap map_out, const __grid_constant__ CUtensorMap map_resid,
                     const GemmParams p) {
  constexpr int CG = 2;
  constexpr uint32_t B_ROWS = BLOCK_N / CG;                      // weight rows this CTA loads per stage
  constexpr uint32_t B_STAGE_BYTES = B_ROWS * GEMM_BLOCK_K * 2;
  constexpr uint32_t SLAB_BYTES = GEMM_AST_MAX_KB * GEMM_A_STAGE_BYTES;
  constexpr uint32_t IDESC = umma_idesc_bf16_f32(GEMM_BLOCK_M * CG, BLOCK_N);
  constexpr uint32_t TMEM_COLS = (2 * BLOCK_N <= 128) ? 128 : ((2 * BLOCK_N <= 256) ? 256 : 512);
  constexpr int NEPI = gemm_epi_warps(BLOCK_N);
  static_assert(BLOCK_N % 64 == 0 && BLOCK_N <= 256, "BLOCK_N must be 64..256 in steps of 64");
  static_assert(B_STAGE_BYTES % 1024 == 0, "weight stages must keep the 1024-byte swizzle alignment");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* ring = smem + SLAB_BYTES;                                        // [nstage] weight half-tiles
  uint8_t* staging = ring + (size_t)p.nstage * B_STAGE_BYTES;               // [NEPI][4 KB], 1024-aligned
  uint64_t* bars = reinterpret_cast<uint64_t*>(staging + (size_t)NEPI * GEMM_STAGING_BYTES);
  uint64_t* b_full = bars;                                 // [nstage]  TMA -> MMA   (leader's copy is the one used)
  uint64_t* b_empty = bars + p.nstage;                     // [nstage]  MMA -> TMA   (one per CTA)
  uint64_t* a_full = bars + 2 * p.nstage;                  // [MAX_KB]  TMA -> MMA   (leader's copy)
  uint64_t* a_empty = a_full + GEMM_AST_MAX_KB;            // [MAX_KB]  MMA -> TMA   (one per CTA)
  uint64_t* acc_full = a_empty + GEMM_AST_MAX_KB;          // [2]       MMA -> epilogue     (one per CTA)
  uint64_t* acc_empty = acc_full + 2;                      // [2]       epilogue -> MMA     (leader's copy)
  uint64_t* resid_bar = acc_empty + 2;                     // [NEPI]    residual box landed (one per epilogue warp)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(resid_bar + NEPI);
  float* bias_smem = reinterpret_cast<float*>(tmem_slot + 4);               // [NEPI][64]

  const int warp = uniform_warp_idx();
  const int lane = threadIdx.x & 31;
  const int rank = (int)cluster_ctarank();
  const bool leader = rank == 0;
  const int n_tiles = p.N / BLOCK_N;
  const int m_blocks = p.M / (GEMM_BLOCK_M * CG);
  const int group = blockIdx.x / CG, n_groups = gridDim.x / CG;
  const int num_k = p.K / GEMM_BLOCK_K;                    // <= GEMM_AST_MAX_KB (checked by the host)

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    tma_prefetch_desc(&map_out);
    if (EPI & EPI_BIAS_RESID) tma_prefetch_desc(&map_resid);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < p.nstage; ++i) {
      mbar_init(&b_full[i], CG);               // one arrival per CTA of the pair
      mbar_init(&b_empty[i], 1);
    }
    for (int i = 0; i < GEMM_AST_MAX_KB; ++i) {
      mbar_init(&a_full[i], CG);
      mbar_init(&a_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], NEPI * CG);
    }
    for (int i = 0; i < NEPI; ++i) mbar_init(&resid_bar[i], 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc_2sm(tmem_slot, TMEM_COLS);
  tc_fence_before();
  cluster_sync_all();                          // the peer's barriers are initialised before anyone signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                                  // the producer of A has finished
  pdl_launch_dependents();

  if (warp == 0) {
    // ===== TMA producer: this CTA's 128 rows of A once per row block, its half of every weight tile; the
    // warp runs the loop converged, one elected lane issues =====
    {
      int s = 0;
      uint32_t ph = 0, mbi = 0;
      for (int mb = group; mb < m_blocks; mb += n_groups, ++mbi) {
        const int m0 = (p.reverse ? m_blocks - 1 - mb : mb) * (GEMM_BLOCK_M * CG) + rank * GEMM_BLOCK_M;
        for (int nt = 0; nt < n_tiles; ++nt) {
          const int n0 = nt * BLOCK_N + rank * (int)B_ROWS;
          for (int kb = 0; kb < num_k; ++kb) {
            if (nt == 0) {
              mbar_wait(&a_empty[kb], (mbi & 1) ^ 1);      // the previous row block is done with this k-block
              if (elect_one()) {
                if (leader) mbar_arrive_expect_tx(&a_full[kb], GEMM_A_STAGE_BYTES * CG);
                else mbar_arrive_leader(&a_full[kb]);
                tma_load_2d_2sm(smem + (size_t)kb * GEMM_A_STAGE_BYTES, &map_a, &a_full[kb], kb * GEMM_BLOCK_K, m0);
              }
              __syncwarp();
            }
            mbar_wait(&b_empty[s], ph ^ 1);
#ifdef B200_GEMM_TIMING
            const bool skip_b = g_gemm_timing[7] == 5 && (kb & 1);   // experiment: half of the weight loads dropped
#else
            const bool skip_b = false;
#endif
            if (elect_one()) {
              if (leader) mbar_arrive_expect_tx(&b_full[s], skip_b ? 0u : B_STAGE_BYTES * CG);
              else mbar_arrive_leader(&b_full[s]);
              if (!skip_b) tma_load_2d_2sm(ring + (size_t)s * B_STAGE_BYTES, &map_b, &b_full[s], kb * GEMM_BLOCK_K, n0);
            }
            __syncwarp();
            if (++s == p.nstage) { s = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the leader CTA's warp 1 runs the loop converged, one elected lane issues =====
    if (leader) {
      int s = 0;
      uint32_t ph = 0, it = 0, mbi = 0;
      const uint32_t slab_addr = smem_u32(smem), ring_addr = smem_u32(ring);
      for (int mb = group; mb < m_blocks; mb += n_groups, ++mbi) {
        for (int nt = 0; nt < n_tiles; ++nt, ++it) {
          const uint32_t as = it & 1, aph = (it >> 1) & 1;
#ifdef B200_GEMM_TIMING
          const long long g0 = clock64();
#endif
          mbar_wait(&acc_empty[as], aph ^ 1);
          tc_fence_after();
#ifdef B200_GEMM_TIMING
          long long g1 = clock64(), gw = 0;
          if (blockIdx.x == 0 && lane == 0) atomicAdd(&g_gemm_timing[0], (unsigned long long)(g1 - g0));
#endif
          const uint32_t tacc = tmem_base + as * BLOCK_N;
          for (int kb = 0; kb < num_k; ++kb) {
#ifdef B200_GEMM_TIMING
            const long long w0 = clock64();
#endif
            if (nt == 0) mbar_wait(&a_full[kb], mbi & 1);
            mbar_wait(&b_full[s], ph);
            tc_fence_after();
#ifdef B200_GEMM_TIMING
            gw += clock64() - w0;
#endif
            const uint64_t da = umma_smem_desc_sw128(slab_addr + (uint32_t)kb * GEMM_A_STAGE_BYTES);
            const uint64_t db = umma_smem_desc_sw128(ring_addr + (uint32_t)s * B_STAGE_BYTES);
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < GEMM_BLOCK_K / 16; ++k)
                umma_bf16_2sm(tacc, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), IDESC, (uint32_t)((kb | k) != 0));
              umma_commit_2sm(&b_empty[s]);
              if (nt == n_tiles - 1) umma_commit_2sm(&a_empty[kb]);   // last use of this k-block of the slab
            }
            __syncwarp();
            if (++s == p.nstage) { s = 0; ph ^= 1; }
          }
          if (elect_one()) umma_commit_2sm(&acc_full[as]);
          __syncwarp();
#ifdef B200_GEMM_TIMING
          if (blockIdx.x == 0 && lane == 0) {
            const long long g2 = clock64();
            atomicAdd(&g_gemm_timing[1], (unsigned long long)gw);
            atomicAdd(&g_gemm_timing[2], (unsigned long long)(g2 - g1 - gw));
            atomicAdd(&g_gemm_timing[3], 1ull);
            atomicAdd(&g_gemm_timing[4], (unsigned long long)(g2 - g0));
          }
#endif
        }
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue warp: TMEM lanes [32 wq, +32) x columns [64 chunk, +64) of this CTA's 128 rows =====
    const int e = warp - 4;
    const int wq = e & 3;         // == warp % 4: the TMEM lane group this warp may read
    const int chunk = e >> 2;
    uint8_t* stg = staging + (size_t)e * GEMM_STAGING_BYTES;
    float* sbias = bias_smem + e * GEMM_OUT_CHUNK;
    uint64_t* rbar = &resid_bar[e];
    uint8_t* my_row = stg + lane * 128;
    const int sw = lane & 7;       // 128-byte swizzle: 16-byte chunk j of row r lives at chunk j ^ (r % 8)
    uint32_t it = 0;
    for (int mb = group; mb < m_blocks; mb += n_groups) {
      for (int nt = 0; nt < n_tiles; ++nt, ++it) {
        const uint32_t as = it & 1, aph = (it >> 1) & 1;
        const int m0 = (p.reverse ? m_blocks - 1 - mb : mb) * (GEMM_BLOCK_M * CG) + rank * GEMM_BLOCK_M + wq * 32;
        const int n0 = nt * BLOCK_N + chunk * GEMM_OUT_CHUNK;
        gemm_epi_tile<EPI, CG>(map_out, map_resid, p, m0, n0, stg, sbias, rbar, my_row, sw, lane,
                               tmem_base + ((uint32_t)(wq * 32) << 16) + as * BLOCK_N + chunk * GEMM_OUT_CHUNK,
                               &acc_full[as], aph, &acc_empty[as], it);
      }
    }
    if (lane == 0) tma_store_wait_all();
  }

  tc_fence_before();
  cluster_sync_all();                          // no CTA leaves while its peer may still signal its barriers
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_2sm(tmem_base, TMEM_COLS);
  }
}

// GEMM + bias + residual + LayerNorm in one kernel, for N == 2 * BLOCK_N (hidden 384: out-proj and FFN2 of
// the MiniLM layer), CTA pairs.  The two TMEM accumulator stages hold the two column tiles of the SAME
// 256-row block, so after the second tile's MMAs a CTA has whole rows on chip:
//   pass 1  (per column tile, as soon as its MMAs are done -- the first overlaps the second tile's MMAs)
//           y = acc + bias + residual, written back to TMEM as fp32; per-row sum and sum of squares
//   stats   the three chunk warps of a TMEM lane group exchange their partials through shared memory
//   pass 2  y re-read from TMEM, the accumulator stage handed back, (y - mean) * rstd * gamma + beta -> bf16
//           -> swizzled staging box -> TMA store
// The separate LayerNorm kernel (read + write of the hidden state, 37 us per call at M = 131072) and the bf16
// round trip of y disappear; statistics are taken on the fp32 values.  Output rows may alias the residual
// rows (x is updated in place): a pair reads all residual tiles of its row block before it stores any.
template <int BLOCK_N>
__global__ void __launch_bounds__(gemm_threads(BLOCK_N), 1)
gemm_ln_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
               const __grid_constant__ CUtensorMap map_out, const __grid_constant__ CUtensorMap map_resid,
               const GemmParams p) {
  constexpr int CG = 2;
  constexpr int N_TOTAL = 2 * BLOCK_N;
  constexpr uint32_t B_ROWS = BLOCK_N / CG;
  constexpr uint32_t B_STAGE_BYTES = B_ROWS * GEMM_BLOCK_K * 2;
  constexpr uint32_t STAGE_BYTES = GEMM_A_STAGE_BYTES + B_STAGE_BYTES;
  constexpr uint32_t IDESC = umma_idesc_bf16_f32(GEMM_BLOCK_M * CG, BLOCK_N);
  constexpr uint32_t TMEM_COLS = 512;
  constexpr int NEPI = gemm_epi_warps(BLOCK_N);
  constexpr int NCHUNK = BLOCK_N / GEMM_OUT_CHUNK;       // chunk warps per TMEM lane group
  static_assert(BLOCK_N == 192, "written for 192-column tiles (two accumulator stages in 512 TMEM columns)");
  static_assert(STAGE_BYTES % 1024 == 0, "stages must keep the 1024-byte swizzle alignment");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* staging = smem + (size_t)p.nstage * STAGE_BYTES;                 // [NEPI][4 KB], 1024-aligned
  uint64_t* bars = reinterpret_cast<uint64_t*>(staging + (size_t)NEPI * GEMM_STAGING_BYTES);
  uint64_t* full = bars;                       // [nstage]  TMA -> MMA          (leader's copy is the one used)
  uint64_t* empty = bars + p.nstage;           // [nstage]  MMA -> TMA          (one per CTA)
  uint64_t* acc_full = bars + 2 * p.nstage;    // [2]       MMA -> epilogue     (one per CTA)
  uint64_t* acc_empty = acc_full + 2;          // [2]       epilogue -> MMA     (leader's copy is the one used)
  uint64_t* resid_bar = acc_empty + 2;         // [NEPI]    residual box landed (one per epilogue warp)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(resid_bar + NEPI);
  float* vecs = reinterpret_cast<float*>(tmem_slot + 4);                    // bias, gamma, beta: [3][N_TOTAL]
  float* xs = vecs + 3 * N_TOTAL;                                           // [2 parities][NCHUNK][128] {sum, sumsq}

  const int warp = uniform_warp_idx();
  const int lane = threadIdx.x & 31;
  const int rank = (int)cluster_ctarank();
  const bool leader = rank == 0;
  const int m_blocks = p.M / (GEMM_BLOCK_M * CG);
  const int group = blockIdx.x / CG, n_groups = gridDim.x / CG;
  const int num_k = p.K / GEMM_BLOCK_K;

  for (int i = threadIdx.x; i < N_TOTAL; i += blockDim.x) {
    vecs[i] = __ldg(p.bias + i);
    vecs[N_TOTAL + i] = __ldg(p.gamma + i);
    vecs[2 * N_TOTAL + i] = __ldg(p.beta + i);
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    tma_prefetch_desc(&map_out);
    tma_prefetch_desc(&map_resid);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < p.nstage; ++i) {
      mbar_init(&full[i], CG);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], NEPI * CG);
    }
    for (int i = 0; i < NEPI; ++i) mbar_init(&resid_bar[i], 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc_2sm(tmem_slot, TMEM_COLS);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                                  // the producers of A and of the residual have finished
  pdl_launch_dependents();

  if (warp == 0) {
    // ===== TMA producer (converged loop, one elected lane issues) =====
    int s = 0;
    uint32_t ph = 0;
    for (int mb = group; mb < m_blocks; mb += n_groups) {
      const int m0 = (p.reverse ? m_blocks - 1 - mb : mb) * (GEMM_BLOCK_M * CG) + rank * GEMM_BLOCK_M;
      for (int nt = 0; nt < 2; ++nt) {
        const int n0 = nt * BLOCK_N + rank * (int)B_ROWS;
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(&empty[s], ph ^ 1);
          uint8_t* sa = smem + (size_t)s * STAGE_BYTES;
          if (elect_one()) {
            if (leader) mbar_arrive_expect_tx(&full[s], STAGE_BYTES * CG);
            else mbar_arrive_leader(&full[s]);
            tma_load_2d_2sm(sa, &map_a, &full[s], kb * GEMM_BLOCK_K, m0);
            tma_load_2d_2sm(sa + GEMM_A_STAGE_BYTES, &map_b, &full[s], kb * GEMM_BLOCK_K, n0);
          }
          __syncwarp();
          if (++s == p.nstage) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA; converged loop, one elected lane issues): tile nt -> accumulator stage nt =====
    if (leader) {
      int s = 0;
      uint32_t ph = 0, mbi = 0;
      const uint32_t smem_addr = smem_u32(smem);
      for (int mb = group; mb < m_blocks; mb += n_groups, ++mbi) {
        for (int nt = 0; nt < 2; ++nt) {
          mbar_wait(&acc_empty[nt], (mbi & 1) ^ 1);
          tc_fence_after();
          const uint32_t tacc = tmem_base + nt * BLOCK_N;
          for (int kb = 0; kb < num_k; ++kb) {
            mbar_wait(&full[s], ph);
            tc_fence_after();
            const uint32_t a_addr = smem_addr + (uint32_t)s * STAGE_BYTES;
            const uint64_t da = umma_smem_desc_sw128(a_addr);
            const uint64_t db = umma_smem_desc_sw128(a_addr + GEMM_A_STAGE_BYTES);
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < GEMM_BLOCK_K / 16; ++k)
                umma_bf16_2sm(tacc, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), IDESC, (uint32_t)((kb | k) != 0));
              umma_commit_2sm(&empty[s]);
            }
            __syncwarp();
            if (++s == p.nstage) { s = 0; ph ^= 1; }
          }
          if (elect_one()) umma_commit_2sm(&acc_full[nt]);
          __syncwarp();
        }
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue warp: TMEM lanes [32 wq, +32) x columns [64 chunk, +64) of BOTH column tiles =====
    const int e = warp - 4;
    const int wq = e & 3;
    const int chunk = e >> 2;
    uint8_t* stg = staging + (size_t)e * GEMM_STAGING_BYTES;
    uint64_t* rbar = &resid_bar[e];
    const uint32_t s_row = smem_u32(stg) + lane * 128;
    const uint32_t s_vecs = smem_u32(vecs), s_xs = smem_u32(xs);
    const int sw = lane & 7;       // 128-byte swizzle: 16-byte chunk j of row r lives at chunk j ^ (r % 8)
    const int row = wq * 32 + lane;
    const uint32_t tlane = tmem_base + ((uint32_t)(wq * 32) << 16) + chunk * GEMM_OUT_CHUNK;
    const float inv_n = 1.0f / (float)N_TOTAL;
    uint32_t mbi = 0, rcount = 0;
    for (int mb = group; mb < m_blocks; mb += n_groups, ++mbi) {
      const int m0 = (p.reverse ? m_blocks - 1 - mb : mb) * (GEMM_BLOCK_M * CG) + rank * GEMM_BLOCK_M + wq * 32;
      float sum = 0.f, sumsq = 0.f;
      // ---- pass 1: y = acc + bias + residual back into TMEM, row statistics ----
#pragma unroll 1
      for (int nt = 0; nt < 2; ++nt, ++rcount) {
        const int n0 = nt * BLOCK_N + chunk * GEMM_OUT_CHUNK;
        __syncwarp();                          // every lane is done reading the box (previous residual tile)
        if (lane == 0) {
          tma_store_wait_read();               // the box is no longer being read by an earlier store
          mbar_arrive_expect_tx(rbar, GEMM_STAGING_BYTES);
          tma_load_2d(stg, &map_resid, rbar, n0, m0);
        }
        __syncwarp();
        mbar_wait(&acc_full[nt], mbi & 1);
        tc_fence_after();
        uint32_t v[64];
        tmem_ld64(tlane + nt * BLOCK_N, v);
        tmem_ld_wait();
        mbar_wait(rbar, rcount & 1);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint4 b0 = lds128(s_vecs + (n0 + j * 8) * 4), b1 = lds128(s_vecs + (n0 + j * 8 + 4) * 4);
          const uint4 r = lds128(s_row + ((j ^ sw) << 4));
          const __nv_bfloat162* r2 = reinterpret_cast<const __nv_bfloat162*>(&r);
          const float bb[8] = {__uint_as_float(b0.x), __uint_as_float(b0.y), __uint_as_float(b0.z), __uint_as_float(b0.w),
                               __uint_as_float(b1.x), __uint_as_float(b1.y), __uint_as_float(b1.z), __uint_as_float(b1.w)};
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float y0 = __uint_as_float(v[j * 8 + 2 * i]) + bb[2 * i] + __low2float(r2[i]);
            const float y1 = __uint_as_float(v[j * 8 + 2 * i + 1]) + bb[2 * i + 1] + __high2float(r2[i]);
            sum += y0 + y1;
            sumsq = fmaf(y0, y0, sumsq);
            sumsq = fmaf(y1, y1, sumsq);
            v[j * 8 + 2 * i] = __float_as_uint(y0);
            v[j * 8 + 2 * i + 1] = __float_as_uint(y1);
          }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint32_t w[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) w[i] = v[q * 16 + i];
          tmem_st16(tlane + nt * BLOCK_N + q * 16, w);
        }
      }
      tmem_st_wait();
      // ---- row statistics across the chunk warps of this lane group ----
      const uint32_t xs_par = s_xs + (mbi & 1) * (NCHUNK * 128 * 8);
      sts32(xs_par + (chunk * 128 + row) * 8, __float_as_uint(sum));
      sts32(xs_par + (chunk * 128 + row) * 8 + 4, __float_as_uint(sumsq));
      named_bar_sync(1 + wq, NCHUNK * 32);
      float tot = 0.f, totsq = 0.f;
#pragma unroll
      for (int c = 0; c < NCHUNK; ++c) {
        tot += __uint_as_float(lds32(xs_par + (c * 128 + row) * 8));
        totsq += __uint_as_float(lds32(xs_par + (c * 128 + row) * 8 + 4));
      }
      const float mean = tot * inv_n;
      const float rstd = rsqrtf(fmaxf(totsq * inv_n - mean * mean, 0.f) + p.eps);
      const float shift = -mean * rstd;
      // ---- pass 2: normalise, affine, bf16, TMA store ----
#pragma unroll 1
      for (int nt = 0; nt < 2; ++nt) {
        const int n0 = nt * BLOCK_N + chunk * GEMM_OUT_CHUNK;
        uint32_t v[64];
        tmem_ld64(tlane + nt * BLOCK_N, v);
        tmem_ld_wait();
        tc_fence_before();         // this warp's share of the accumulator stage is in registers
        __syncwarp();
        if (lane == 0) {
          mbar_arrive_leader(&acc_empty[nt]);
          tma_store_wait_read();   // the previous store has finished reading this warp's box
        }
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint4 g0 = lds128(s_vecs + (N_TOTAL + n0 + j * 8) * 4), g1 = lds128(s_vecs + (N_TOTAL + n0 + j * 8 + 4) * 4);
          const uint4 e0 = lds128(s_vecs + (2 * N_TOTAL + n0 + j * 8) * 4), e1 = lds128(s_vecs + (2 * N_TOTAL + n0 + j * 8 + 4) * 4);
          const float gg[8] = {__uint_as_float(g0.x), __uint_as_float(g0.y), __uint_as_float(g0.z), __uint_as_float(g0.w),
                               __uint_as_float(g1.x), __uint_as_float(g1.y), __uint_as_float(g1.z), __uint_as_float(g1.w)};
          const float ee[8] = {__uint_as_float(e0.x), __uint_as_float(e0.y), __uint_as_float(e0.z), __uint_as_float(e0.w),
                               __uint_as_float(e1.x), __uint_as_float(e1.y), __uint_as_float(e1.z), __uint_as_float(e1.w)};
          float f[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] = fmaf(fmaf(__uint_as_float(v[j * 8 + i]), rstd, shift), gg[i], ee[i]);
          uint4 o;
          __nv_bfloat162 t0 = __floats2bfloat162_rn(f[0], f[1]);
          __nv_bfloat162 t1 = __floats2bfloat162_rn(f[2], f[3]);
          __nv_bfloat162 t2 = __floats2bfloat162_rn(f[4], f[5]);
          __nv_bfloat162 t3 = __floats2bfloat162_rn(f[6], f[7]);
          o.x = *reinterpret_cast<uint32_t*>(&t0); o.y = *reinterpret_cast<uint32_t*>(&t1);
          o.z = *reinterpret_cast<uint32_t*>(&t2); o.w = *reinterpret_cast<uint32_t*>(&t3);
          sts128(s_row + ((j ^ sw) << 4), o);
        }
        fence_proxy_async();       // generic-proxy writes -> visible to the TMA (async proxy)
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&map_out, stg, n0, m0);
          tma_store_commit();
        }
      }
    }
    if (lane == 0) tma_store_wait_all();
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_2sm(tmem_base, TMEM_COLS);
  }
}

// =============================================================================================
// K3: attention for one (sequence, head) per CTA.  8 warps; K and V of the head sit in shared
// memory (row-major, rows padded by 16 bytes so that ldmatrix phases are bank-conflict free); each
// warp owns 16 query rows per pass and walks the keys in blocks of 64 with an online softmax (fp32);
// QK^T and PV run on mma.sync m16n8k16 bf16 tiles whose B fragments come from ldmatrix (.trans for
// V).  qkv is [tokens, 3H] with columns [Q | K | V], head h at offset h*DH inside each third; ctx
// is [tokens, H].  Query tiles that lie entirely in the padding are skipped: nothing reads them
// (padded keys are masked, padded rows are excluded from pooling).
// =============================================================================================
B200_DEVINL void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
B200_DEVINL void ldmatrix_x4(uint32_t (&r)[4], const void* smem_row) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_u32(smem_row)));
}
B200_DEVINL void ldmatrix_x4_trans(uint32_t (&r)[4], const void* smem_row) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_u32(smem_row)));
}
// 2^x on the SFU, flush-to-zero; 2^(-inf) = 0
B200_DEVINL float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
B200_DEVINL uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}

template <int DH>
__global__ void __launch_bounds__(256, (DH == 32) ? 2 : 1)
attention_kernel(const __nv_bfloat16* __restrict__ qkv, const int32_t* __restrict__ lens, int S, int H,
                 __nv_bfloat16* __restrict__ ctx) {
  constexpr int KSTR = DH + 8;          // row stride (elements) of the K and V tiles
  constexpr int KSTEPS = DH / 16;       // k-steps of QK^T
  constexpr int ONT = DH / 8;           // output n-tiles
  extern __shared__ __align__(16) uint8_t att_smem[];
  __nv_bfloat16* Ks = reinterpret_cast<__nv_bfloat16*>(att_smem);   // [S][KSTR]
  __nv_bfloat16* Vs = Ks + (size_t)S * KSTR;                         // [S][KSTR]

  const int h = blockIdx.x, b = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int len = max(1, min(lens[b], S));
  const int len16 = (len + 15) & ~15;
  const size_t tok0 = (size_t)b * S;
  const int ld = 3 * H;

  // stage K and V for keys [0, len16); keys >= len are zero-filled (and masked below)
  constexpr int PIECES = DH / 8;  // 16-byte pieces per row
  for (int i = threadIdx.x; i < len16 * PIECES; i += blockDim.x) {
    const int t = i / PIECES, pc = i % PIECES;
    uint4 kv = make_uint4(0, 0, 0, 0), vv = make_uint4(0, 0, 0, 0);
    if (t < len) {
      const __nv_bfloat16* base = qkv + (tok0 + t) * ld + h * DH + pc * 8;
      kv = *reinterpret_cast<const uint4*>(base + H);
      vv = *reinterpret_cast<const uint4*>(base + 2 * H);
    }
    *reinterpret_cast<uint4*>(Ks + (size_t)t * KSTR + pc * 8) = kv;
    *reinterpret_cast<uint4*>(Vs + (size_t)t * KSTR + pc * 8) = vv;
  }
  __syncthreads();

  const float scale_log2 = 1.0f;   // log2(e) / sqrt(DH) is folded into the query projection (encoder.cu)
  const int r = lane >> 2, cq = (lane & 3) * 2;
  const int lm = lane >> 3, lr = lane & 7;   // ldmatrix: this lane addresses row lr of matrix lm

  for (int q0 = warp * 16; q0 < len16; q0 += 8 * 16) {
    // Q fragments for this warp's 16 rows
    uint32_t qa[KSTEPS][4];
    {
      const __nv_bfloat16* q_lo = qkv + (tok0 + q0 + r) * ld + h * DH;
      const __nv_bfloat16* q_hi = q_lo + (size_t)8 * ld;
#pragma unroll
      for (int ks = 0; ks < KSTEPS; ++ks) {
        qa[ks][0] = *reinterpret_cast<const uint32_t*>(q_lo + ks * 16 + cq);
        qa[ks][1] = *reinterpret_cast<const uint32_t*>(q_hi + ks * 16 + cq);
        qa[ks][2] = *reinterpret_cast<const uint32_t*>(q_lo + ks * 16 + 8 + cq);
        qa[ks][3] = *reinterpret_cast<const uint32_t*>(q_hi + ks * 16 + 8 + cq);
      }
    }
    float o[ONT][4];
#pragma unroll
    for (int i = 0; i < ONT; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m_lo = -INFINITY, m_hi = -INFINITY, l_lo = 0.f, l_hi = 0.f;

    for (int kb = 0; kb < len16; kb += 64) {
      const int nkt = min(8, (len16 - kb) >> 3);  // 8-key tiles in this block (even: len16 % 16 == 0)
      float s[8][4];
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
        if (nt < nkt) {
          // matrices of one ldmatrix.x4: K[keys nt*8.., dims 8*lm .. 8*lm+7] -> b0/b1 of two k-steps
          const __nv_bfloat16* krow = Ks + (size_t)(kb + nt * 8 + lr) * KSTR + lm * 8;
#pragma unroll
          for (int kk = 0; kk < KSTEPS / 2; ++kk) {
            uint32_t kf[4];
            ldmatrix_x4(kf, krow + kk * 32);
            mma_bf16_16816(s[nt], qa[2 * kk], kf[0], kf[1]);
            mma_bf16_16816(s[nt], qa[2 * kk + 1], kf[2], kf[3]);
          }
        }
      }
      // block row max on the raw scores; keys >= len are masked only in the block that holds the
      // boundary (full blocks skip the compare/select per element)
      if (kb + 64 > len) {
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          const int key = kb + nt * 8 + cq;
          const bool v0 = (nt < nkt) && key < len, v1 = (nt < nkt) && key + 1 < len;
          s[nt][0] = v0 ? s[nt][0] : -INFINITY;
          s[nt][1] = v1 ? s[nt][1] : -INFINITY;
          s[nt][2] = v0 ? s[nt][2] : -INFINITY;
          s[nt][3] = v1 ? s[nt][3] : -INFINITY;
        }
      }
      float bm_lo = -INFINITY, bm_hi = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        bm_lo = fmaxf(bm_lo, fmaxf(s[nt][0], s[nt][1]));
        bm_hi = fmaxf(bm_hi, fmaxf(s[nt][2], s[nt][3]));
      }
      bm_lo = fmaxf(bm_lo, __shfl_xor_sync(0xffffffffu, bm_lo, 1));
      bm_lo = fmaxf(bm_lo, __shfl_xor_sync(0xffffffffu, bm_lo, 2));
      bm_hi = fmaxf(bm_hi, __shfl_xor_sync(0xffffffffu, bm_hi, 1));
      bm_hi = fmaxf(bm_hi, __shfl_xor_sync(0xffffffffu, bm_hi, 2));
      const float mn_lo = fmaxf(m_lo, bm_lo), mn_hi = fmaxf(m_hi, bm_hi);  // finite: key 0 is always valid
      const float al_lo = fast_exp2((m_lo - mn_lo) * scale_log2), al_hi = fast_exp2((m_hi - mn_hi) * scale_log2);
      m_lo = mn_lo; m_hi = mn_hi;
      l_lo *= al_lo; l_hi *= al_hi;
#pragma unroll
      for (int i = 0; i < ONT; ++i) { o[i][0] *= al_lo; o[i][1] *= al_lo; o[i][2] *= al_hi; o[i][3] *= al_hi; }
      // P = 2^((s - m) * scale * log2 e): one FFMA + one MUFU per element; P is rounded to bf16 for
      // the PV tensor tiles, the row sums accumulate the unrounded values
      const float ms_lo = m_lo * scale_log2, ms_hi = m_hi * scale_log2;
      uint32_t pa[4][4];
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        const float p0 = fast_exp2(fmaf(s[nt][0], scale_log2, -ms_lo)), p1 = fast_exp2(fmaf(s[nt][1], scale_log2, -ms_lo));
        const float p2 = fast_exp2(fmaf(s[nt][2], scale_log2, -ms_hi)), p3 = fast_exp2(fmaf(s[nt][3], scale_log2, -ms_hi));
        l_lo += p0 + p1;
        l_hi += p2 + p3;
        pa[nt >> 1][(nt & 1) * 2 + 0] = pack_bf16(p0, p1);
        pa[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16(p2, p3);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (j * 2 < nkt) {
          // matrices of one ldmatrix.x4.trans: V[keys 16j + 8*(lm&1).., dims 8*(2*ip + (lm>>1))..]
          const __nv_bfloat16* vrow = Vs + (size_t)(kb + j * 16 + (lm & 1) * 8 + lr) * KSTR + (lm >> 1) * 8;
#pragma unroll
          for (int ip = 0; ip < ONT / 2; ++ip) {
            uint32_t vf[4];
            ldmatrix_x4_trans(vf, vrow + ip * 16);
            mma_bf16_16816(o[2 * ip], pa[j], vf[0], vf[1]);
            mma_bf16_16816(o[2 * ip + 1], pa[j], vf[2], vf[3]);
          }
        }
      }
    }
    l_lo += __shfl_xor_sync(0xffffffffu, l_lo, 1);
    l_lo += __shfl_xor_sync(0xffffffffu, l_lo, 2);
    l_hi += __shfl_xor_sync(0xffffffffu, l_hi, 1);
    l_hi += __shfl_xor_sync(0xffffffffu, l_hi, 2);
    const float inv_lo = 1.0f / l_lo, inv_hi = 1.0f / l_hi;
    __nv_bfloat16* c_lo = ctx + (tok0 + q0 + r) * H + h * DH + cq;
    __nv_bfloat16* c_hi = c_lo + (size_t)8 * H;
#pragma unroll
    for (int i = 0; i < ONT; ++i) {
      *reinterpret_cast<uint32_t*>(c_lo + i * 8) = pack_bf16(o[i][0] * inv_lo, o[i][1] * inv_lo);
      *reinterpret_cast<uint32_t*>(c_hi + i * 8) = pack_bf16(o[i][2] * inv_hi, o[i][3] * inv_hi);
    }
  }
}

// =============================================================================================
// K7: masked mean over the valid tokens, then x / max(||x||, 1e-12).  One CTA per sequence.  A thread
// owns 8 consecutive features (one 16-byte load per token row) of every ng-th token, ng = 256 / (H / 8)
// token rows in flight per pass; the ng partial sums per feature meet in shared memory and are added
// in a fixed order (run-to-run bit-reproducible).  H % 8 == 0, H <= 1024.
// =============================================================================================
__global__ void __launch_bounds__(256) pool_normalize_kernel(const __nv_bfloat16* __restrict__ x,
                                                             const int32_t* __restrict__ lens, int S, int H,
                                                             float* __restrict__ out) {
  __shared__ float part[2048];                 // [ng][H], ng * H <= 256 * 8
  __shared__ float red[8];
  const int b = blockIdx.x;
  const int len = max(0, min(lens[b], S));
  const float denom = fmaxf((float)len, 1e-9f);
  const int nf = H >> 3, ng = max(1, (int)blockDim.x / nf);
  const int g = threadIdx.x / nf, f = threadIdx.x - g * nf;
  if (g < ng) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    const uint4* base = reinterpret_cast<const uint4*>(x + (size_t)b * S * H) + f;
    const int row_stride = H >> 3;               // uint4 per token row
    int t = g;
    for (; t + ng < len; t += 2 * ng) {          // two independent loads in flight
      const uint4 u0 = __ldg(base + (size_t)t * row_stride), u1 = __ldg(base + (size_t)(t + ng) * row_stride);
      const uint32_t w0[4] = {u0.x, u0.y, u0.z, u0.w}, w1[4] = {u1.x, u1.y, u1.z, u1.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        acc[2 * i] += __uint_as_float(w0[i] << 16) + __uint_as_float(w1[i] << 16);
        acc[2 * i + 1] += __uint_as_float(w0[i] & 0xffff0000u) + __uint_as_float(w1[i] & 0xffff0000u);
      }
    }
    if (t < len) {
      const uint4 u0 = __ldg(base + (size_t)t * row_stride);
      const uint32_t w0[4] = {u0.x, u0.y, u0.z, u0.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        acc[2 * i] += __uint_as_float(w0[i] << 16);
        acc[2 * i + 1] += __uint_as_float(w0[i] & 0xffff0000u);
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) part[g * H + f * 8 + i] = acc[i];
  }
  __syncthreads();
  float v[4] = {0.f, 0.f, 0.f, 0.f};           // H <= 1024 with 256 threads
  float ss = 0.f;
  for (int k = 0, c = threadIdx.x; c < H; c += blockDim.x, ++k) {
    float a = 0.f;
    for (int gg = 0; gg < ng; ++gg) a += part[gg * H + c];
    a /= denom;
    v[k] = a;
    ss = fmaf(a, a, ss);
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
  for (int i = 0; i < (int)(blockDim.x >> 5); ++i) tot += red[i];
  const float inv = 1.0f / fmaxf(sqrtf(tot), 1e-12f);
  for (int k = 0, c = threadIdx.x; c < H; c += blockDim.x, ++k) out[(size_t)b * H + c] = v[k] * inv;
}

}  // namespace b200
