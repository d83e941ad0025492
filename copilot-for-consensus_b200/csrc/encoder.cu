// encoder.cu -- host side of the BERT-class encoder behind the C ABI (include/b200_retrieval.h).
//
// Replaces the arithmetic behind SentenceTransformerEmbeddingProvider.embed ->
// self.model.encode(...) (adapters/copilot_embedding/copilot_embedding/sentence_transformer_provider.py:93)
// for BERT-architecture models: token ids in, L2-normalised sentence embeddings out.  Tokenisation
// stays on the host, in Python.
//
// Data layout in HBM
//   weights  bf16, nn.Linear layout [out, in] (= K-major B operand for tcgen05); Q/K/V packed as
//            one [3H, H]; embedding tables bf16; biases and LayerNorm parameters fp32
//   x        bf16 [M, H]   hidden state            qkv  bf16 [M, 3H]
//   ctx      bf16 [M, H]   attention output        y    bf16 [M, H]  pre-LayerNorm sum
//   ffn      bf16 [M, I]   GELU(W1 x + b1)         M = batch * seq rounded up to 128 rows
#include <map>
#include <string>
#include <vector>

#include "attention_tc.cuh"
#include "attention_tc64.cuh"
#include "ffn_fused.cuh"
#include "encoder_kernels.cuh"
#include "host_util.h"

using namespace b200;

namespace {

struct LayerWeights {
  DevBuf wqkv, wo, w1, w2;              // bf16
  DevBuf bqkv, bo, b1, b2;              // fp32
  DevBuf ln1_g, ln1_b, ln2_g, ln2_b;    // fp32
  CUtensorMap map_wqkv, map_wo, map_w1, map_w2;           // box rows = BLOCK_N
  CUtensorMap map_wqkv_h, map_wo_h, map_w1_h, map_w2_h;   // box rows = BLOCK_N / 2 (one CTA's share in a CTA pair)
  CUtensorMap map_w1_f;                                   // box rows = 64: one CTA's share of a 128-column FFN slice (ffn_fused_kernel)
  // deferred LayerNorm (EPI_ROWFOLD): weights with the preceding LayerNorm's gamma folded into their columns,
  // their row sums c[n] (of the bf16 values), and biases with W beta added
  DevBuf wqkv_g, w1_g, cqkv, c1, bqkv_g, b1_g;
  CUtensorMap map_wqkv_gh, map_w1_gh;                     // box rows = BLOCK_N / 2
};

uint16_t f32_to_bf16_host(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

int upload_bf16(DevBuf& dst, const float* src, size_t n) {
  std::vector<uint16_t> tmp(n);
  for (size_t i = 0; i < n; ++i) tmp[i] = f32_to_bf16_host(src[i]);
  B200_TRY(dst.alloc(n * 2));
  B200_CUDA(cudaMemcpy(dst.p, tmp.data(), n * 2, cudaMemcpyHostToDevice));
  return B200_OK;
}
int upload_f32(DevBuf& dst, const float* src, size_t n) {
  B200_TRY(dst.alloc(n * 4));
  B200_CUDA(cudaMemcpy(dst.p, src, n * 4, cudaMemcpyHostToDevice));
  return B200_OK;
}

}  // namespace

struct b200_encoder {
  b200_encoder_config_t cfg{};
  int device = 0, num_sms = 148;
  bool finalized = false;
  std::map<std::string, std::vector<float>> pending;  // tensors handed over before finalize
  DevBuf word, pos, type0, emb_g, emb_b;
  std::vector<LayerWeights*> layers;
  // activations, sized for m_cap rows
  int64_t m_cap = 0;
  DevBuf x, qkv, ctx, y, ffn, ids, lens, out;
  DevBuf att_kmax2, att_flag;   // attention_tc64: max |k|^2 per (sequence, head); "take the exact-maximum script" flag
  bool defer_capable = false;   // the folded weights of the deferred-LayerNorm path exist (finalize)
  DevBuf ln_ones, ln_zeros;     // [hidden] gamma = 1 / beta = 0 for layer 0 of the deferred-LayerNorm path
  DevBuf ln_stats0, ln_stats1;  // deferred LayerNorm: [m_cap][hidden / 64][2] (sum, sum of squares) of x-buffer / y-buffer rows
  CUtensorMap map_y;                                        // y as an A operand (deferred LayerNorm: FFN1 reads it)
  CUtensorMap map_x, map_ctx, map_ffn;                      // A operands: box {64, 128}
  CUtensorMap map_qkv_st, map_y_st, map_ffn_st, map_x_res;  // epilogue boxes {64, 32}: stores and the residual load
  StreamOrder order;   // calls arriving on different streams share the activation buffers: see host_util.h
  std::mutex mu;
  ~b200_encoder() {
    for (auto* l : layers) delete l;
  }
};

static constexpr int kGemmBlockN = 192;

static int pick_block_n(int N) {
  if (N % 192 == 0) return 192;
  if (N % 128 == 0) return 128;
  if (N % 64 == 0) return 64;
  return 0;
}

struct GemmIO {
  const CUtensorMap* a;      // activations [M, K], box {64, 128}
  const CUtensorMap* w;      // weights [N, K], box {64, BLOCK_N}
  const CUtensorMap* w2;     // weights [N, K], box {64, BLOCK_N / 2} (CTA-pair kernel) or nullptr
  const CUtensorMap* out;    // output [M, N], box {64, 32} (TMA store)
  const CUtensorMap* resid;  // residual [M, N], box {64, 32} (TMA load) or the same as out when unused
};

// Launch attributes of the hot-path kernels: cluster of two (CTA pairs) and, with option pdl=1, programmatic
// stream serialisation (see pdl_wait in common.cuh).  Off by default: measured +0.3 % on the 512 x 256
// batch (3.975 -> 3.960 ms per step) -- the kernels are long and persistent, so their boundaries cost
// little -- which is not worth relaxing the stream order for.  Returns the attribute count.
static int hot_launch_attrs(cudaLaunchAttribute* attr, bool pair) {
  int n = 0;
  if (pair) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = 2; attr[n].val.clusterDim.y = 1; attr[n].val.clusterDim.z = 1;
    ++n;
  }
  const bool pdl = opt(OPT_PDL) != 0;
  if (pdl && !prof_enabled()) {     // per-kernel timing records events between launches: keep those runs serial
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  return n;
}

template <int BLOCK_N, int EPI, int CG>
static int launch_gemm_t(const b200_encoder* e, const GemmIO& io, GemmParams p, cudaStream_t st) {
  constexpr int NEPI = gemm_epi_warps(BLOCK_N);
  const size_t stage = (size_t)GEMM_A_STAGE_BYTES + (size_t)(BLOCK_N / CG) * GEMM_BLOCK_K * 2;
  const size_t max_smem = 232448;
  // staging boxes, barriers (2*nstage + 4 + NEPI, at most 8 stages), TMEM slot, per-warp bias slices
  const size_t tail = (size_t)NEPI * GEMM_STAGING_BYTES + 8 * (2 * 8 + 4 + NEPI) + 16 +
                      (size_t)NEPI * GEMM_OUT_CHUNK * 4 * gemm_epi_consts(EPI);
  int nstage = (int)std::min<size_t>(8, (max_smem - 1024 - tail) / stage);
  if (nstage < 2) B200_FAIL(B200_ERR_UNSUPPORTED, "GEMM tile does not fit in shared memory");
  p.nstage = nstage;
  const size_t smem = 1024 + (size_t)nstage * stage + tail;
  auto kern = gemm_bf16_kernel<BLOCK_N, EPI, CG>;
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int tiles = (p.M / (GEMM_BLOCK_M * CG)) * (p.N / BLOCK_N);
  const int grid = std::min(tiles, e->num_sms / CG) * CG;
  {
    B200_PROF((EPI & EPI_BIAS_RESID) ? "gemm_bias_resid" : ((EPI & EPI_BIAS_GELU) ? "gemm_bias_gelu" : "gemm_bias"),
              2.0 * p.M * (double)p.N * p.K,
              2.0 * ((double)p.M * p.K + (double)p.N * p.K + (double)p.M * p.N * ((EPI & EPI_BIAS_RESID) ? 2 : 1)), st);
    if (CG == 1) {
      kern<<<grid, gemm_threads(BLOCK_N), smem, st>>>(*io.a, *io.w, *io.out, *io.resid, p);
    } else {
      cudaLaunchConfig_t cfg{};
      cfg.gridDim = dim3(grid);
      cfg.blockDim = dim3(gemm_threads(BLOCK_N));
      cfg.dynamicSmemBytes = smem;
      cfg.stream = st;
      cudaLaunchAttribute attr[2];
      cfg.attrs = attr; cfg.numAttrs = hot_launch_attrs(attr, true);
      B200_CUDA(cudaLaunchKernelEx(&cfg, kern, *io.a, *io.w2, *io.out, *io.resid, p));
    }
  }
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// A-stationary CTA-pair GEMM (gemm_bf16_ast_kernel): K <= 384, whole 256-row blocks.
template <int BLOCK_N, int EPI>
static int launch_gemm_ast(const b200_encoder* e, const GemmIO& io, GemmParams p, cudaStream_t st) {
  constexpr int NEPI = gemm_epi_warps(BLOCK_N);
  const size_t slab = (size_t)GEMM_AST_MAX_KB * GEMM_A_STAGE_BYTES;
  const size_t stage = (size_t)(BLOCK_N / 2) * GEMM_BLOCK_K * 2;
  const size_t max_smem = 232448;
  // staging boxes, barriers (2*nstage + 2*MAX_KB + 4 + NEPI, at most 8 stages), TMEM slot, per-warp bias slices
  const size_t tail = (size_t)NEPI * GEMM_STAGING_BYTES + 8 * (2 * 8 + 2 * GEMM_AST_MAX_KB + 4 + NEPI) + 16 +
                      (size_t)NEPI * GEMM_OUT_CHUNK * 4;
  const int nstage = (int)std::min<size_t>(8, (max_smem - 1024 - slab - tail) / stage);
  if (nstage < 2) B200_FAIL(B200_ERR_UNSUPPORTED, "A-stationary GEMM tile does not fit in shared memory");
  p.nstage = nstage;
  const size_t smem = 1024 + slab + (size_t)nstage * stage + tail;
  auto kern = gemm_bf16_ast_kernel<BLOCK_N, EPI>;
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int m_blocks = p.M / (2 * GEMM_BLOCK_M);
  const int grid = std::min(m_blocks, e->num_sms / 2) * 2;
  {
    B200_PROF(EPI == EPI_BIAS ? "gemm_bias" : (EPI == EPI_BIAS_GELU ? "gemm_bias_gelu" : "gemm_bias_resid"),
              2.0 * p.M * (double)p.N * p.K,
              2.0 * ((double)p.M * p.K + (double)p.N * p.K + (double)p.M * p.N * (EPI == EPI_BIAS_RESID ? 2 : 1)), st);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(gemm_threads(BLOCK_N));
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    cfg.attrs = attr; cfg.numAttrs = hot_launch_attrs(attr, true);
    B200_CUDA(cudaLaunchKernelEx(&cfg, kern, *io.a, *io.w2, *io.out, *io.resid, p));
  }
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// GEMM + bias + residual + LayerNorm (gemm_ln_kernel): N == 384, whole 256-row blocks.  `io.out` is the
// store map of the normalised output (it may be the residual's own tensor: updated in place).
static int launch_gemm_ln(const b200_encoder* e, const GemmIO& io, GemmParams p, cudaStream_t st) {
  constexpr int BLOCK_N = 192;
  constexpr int NEPI = gemm_epi_warps(BLOCK_N);
  const size_t stage = (size_t)GEMM_A_STAGE_BYTES + (size_t)(BLOCK_N / 2) * GEMM_BLOCK_K * 2;
  const size_t max_smem = 232448;
  // staging boxes, barriers (2*nstage + 4 + NEPI, at most 8 stages), TMEM slot, bias/gamma/beta, row statistics
  const size_t tail = (size_t)NEPI * GEMM_STAGING_BYTES + 8 * (2 * 8 + 4 + NEPI) + 16 + 3 * 2 * BLOCK_N * 4 +
                      2 * (BLOCK_N / GEMM_OUT_CHUNK) * 128 * 8;
  const int nstage = (int)std::min<size_t>(8, (max_smem - 1024 - tail) / stage);
  if (nstage < 2) B200_FAIL(B200_ERR_UNSUPPORTED, "fused GEMM + LayerNorm tile does not fit in shared memory");
  p.nstage = nstage;
  const size_t smem = 1024 + (size_t)nstage * stage + tail;
  auto kern = gemm_ln_kernel<BLOCK_N>;
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int m_blocks = p.M / (2 * GEMM_BLOCK_M);
  const int grid = std::min(m_blocks, e->num_sms / 2) * 2;
  {
    B200_PROF("gemm_bias_resid_ln", 2.0 * p.M * (double)p.N * p.K,
              2.0 * ((double)p.M * p.K + (double)p.N * p.K + (double)p.M * p.N * 2), st);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(gemm_threads(BLOCK_N));
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    cfg.attrs = attr; cfg.numAttrs = hot_launch_attrs(attr, true);
    B200_CUDA(cudaLaunchKernelEx(&cfg, kern, *io.a, *io.w2, *io.out, *io.resid, p));
  }
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// The whole feed-forward block of a layer in one kernel (ffn_fused_kernel): hidden 384, whole 256-row blocks.
static int launch_ffn_fused(const b200_encoder* e, const LayerWeights* L, int M, int I, int reverse, cudaStream_t st) {
  auto kern = ffn_fused_kernel;
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FFN_SMEM_BYTES));
  FfnParams p{};
  p.M = M; p.I = I;
  p.b1 = L->b1.as<float>(); p.b2 = L->b2.as<float>();
  p.gamma = L->ln2_g.as<float>(); p.beta = L->ln2_b.as<float>(); p.eps = e->cfg.ln_eps;
  p.reverse = reverse;
  const int m_blocks = M / (2 * GEMM_BLOCK_M);
  const int grid = std::min(m_blocks, e->num_sms / 2) * 2;
  {
    const double H = FFN_H;
    B200_PROF("ffn_fused", 4.0 * M * H * I, 2.0 * (2.0 * M * H + 2.0 * H * I), st);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(FFN_THREADS);
    cfg.dynamicSmemBytes = FFN_SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    cfg.attrs = attr; cfg.numAttrs = hot_launch_attrs(attr, true);
    B200_CUDA(cudaLaunchKernelEx(&cfg, kern, e->map_x, L->map_w1_f, L->map_w2_h, e->map_x_res, p));
  }
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

template <int EPI>
static int launch_gemm(const b200_encoder* e, const GemmIO& io, const GemmParams& p, cudaStream_t st) {
  // CTA pairs need whole 256-row tiles; option gemm_pair=0 forces single-CTA tiles (A/B timing)
  const bool pair = (p.M % (2 * GEMM_BLOCK_M) == 0) && io.w2 != nullptr && opt(OPT_GEMM_PAIR) != 0;
  // K <= 384: keep the activation slab in shared memory across column tiles (option gemm_ast=0 disables)
  if (pair && pick_block_n(p.N) == 192 && p.K <= GEMM_AST_MAX_KB * GEMM_BLOCK_K && p.N / 192 >= 2 &&
      opt(OPT_GEMM_AST) != 0)
    return launch_gemm_ast<192, EPI>(e, io, p, st);
  switch (pick_block_n(p.N)) {
    case 192: return pair ? launch_gemm_t<192, EPI, 2>(e, io, p, st) : launch_gemm_t<192, EPI, 1>(e, io, p, st);
    case 128: return pair ? launch_gemm_t<128, EPI, 2>(e, io, p, st) : launch_gemm_t<128, EPI, 1>(e, io, p, st);
    case 64: return launch_gemm_t<64, EPI, 1>(e, io, p, st);
  }
  B200_FAIL(B200_ERR_UNSUPPORTED, "GEMM N=%d is not a multiple of 64", p.N);
}

static int weight_map(CUtensorMap* m, const DevBuf& w, int N, int K, int cg = 1) {
  return make_tma_2d_bf16(m, w.p, (uint64_t)N, (uint64_t)K, (uint32_t)(pick_block_n(N) / cg), GEMM_BLOCK_K);
}

static int ensure_activations(b200_encoder* e, int64_t m_rows, int batch, cudaStream_t st) {
  const int H = e->cfg.hidden, I = e->cfg.intermediate;
  B200_TRY(e->lens.grow((size_t)std::max(batch, 1) * 4, 0, st));
  B200_TRY(e->out.grow((size_t)std::max(batch, 1) * H * 4, 0, st));
  B200_TRY(e->att_kmax2.grow((size_t)2 * std::max(batch, 1) * e->cfg.heads * 4, 0, st));   // max |k|^2, then max |q|^2
  B200_TRY(e->att_flag.grow(64, 0, st));
  if (m_rows <= e->m_cap) return B200_OK;
  const int64_t m = (m_rows + 255) / 256 * 256;   // whole 256-row tiles for the CTA-pair GEMM
  B200_TRY(e->ids.grow((size_t)m * 4, 0, st));
  B200_TRY(e->x.grow((size_t)m * H * 2, 0, st));
  B200_TRY(e->qkv.grow((size_t)m * 3 * H * 2, 0, st));
  B200_TRY(e->ctx.grow((size_t)m * H * 2, 0, st));
  B200_TRY(e->y.grow((size_t)m * H * 2, 0, st));
  B200_TRY(e->ffn.grow((size_t)m * I * 2, 0, st));
  // rows beyond the live batch are computed along with it (whole 128-row tiles); keep them finite
  B200_CUDA(cudaMemsetAsync(e->x.p, 0, (size_t)m * H * 2, st));
  B200_CUDA(cudaMemsetAsync(e->ctx.p, 0, (size_t)m * H * 2, st));
  B200_CUDA(cudaMemsetAsync(e->ffn.p, 0, (size_t)m * I * 2, st));
  B200_CUDA(cudaMemsetAsync(e->qkv.p, 0, (size_t)m * 3 * H * 2, st));
  B200_TRY(make_tma_2d_bf16(&e->map_x, e->x.p, (uint64_t)m, (uint64_t)H, GEMM_BLOCK_M, GEMM_BLOCK_K));
  B200_TRY(make_tma_2d_bf16(&e->map_ctx, e->ctx.p, (uint64_t)m, (uint64_t)H, GEMM_BLOCK_M, GEMM_BLOCK_K));
  B200_TRY(make_tma_2d_bf16(&e->map_y, e->y.p, (uint64_t)m, (uint64_t)H, GEMM_BLOCK_M, GEMM_BLOCK_K));
  if (e->defer_capable) {
    const size_t sb = (size_t)m * (H / GEMM_OUT_CHUNK) * 2 * 4;
    B200_TRY(e->ln_stats0.grow(sb, 0, st));
    B200_TRY(e->ln_stats1.grow(sb, 0, st));
    B200_CUDA(cudaMemsetAsync(e->ln_stats0.p, 0, sb, st));
    B200_CUDA(cudaMemsetAsync(e->ln_stats1.p, 0, sb, st));
    B200_CUDA(cudaMemsetAsync(e->y.p, 0, (size_t)m * H * 2, st));
  }
  B200_TRY(make_tma_2d_bf16(&e->map_ffn, e->ffn.p, (uint64_t)m, (uint64_t)I, GEMM_BLOCK_M, GEMM_BLOCK_K));
  B200_TRY(make_tma_2d_bf16(&e->map_qkv_st, e->qkv.p, (uint64_t)m, (uint64_t)(3 * H), 32, GEMM_OUT_CHUNK));
  B200_TRY(make_tma_2d_bf16(&e->map_y_st, e->y.p, (uint64_t)m, (uint64_t)H, 32, GEMM_OUT_CHUNK));
  B200_TRY(make_tma_2d_bf16(&e->map_ffn_st, e->ffn.p, (uint64_t)m, (uint64_t)I, 32, GEMM_OUT_CHUNK));
  B200_TRY(make_tma_2d_bf16(&e->map_x_res, e->x.p, (uint64_t)m, (uint64_t)H, 32, GEMM_OUT_CHUNK));
  e->m_cap = m;
  return B200_OK;
}

static int take(b200_encoder* e, const std::string& name, size_t numel, const std::vector<float>** out) {
  auto it = e->pending.find(name);
  if (it == e->pending.end()) B200_FAIL(B200_ERR_INVALID, "missing tensor '%s'", name.c_str());
  if (it->second.size() != numel)
    B200_FAIL(B200_ERR_DIMENSION, "tensor '%s' has %zu elements, expected %zu", name.c_str(), it->second.size(), numel);
  *out = &it->second;
  return B200_OK;
}

template <int NV>
static int run_encoder(b200_encoder* e, const int32_t* d_ids, const int32_t* d_lens, int batch, int seq, float* d_out,
                       cudaStream_t st) {
  const int H = e->cfg.hidden, I = e->cfg.intermediate, heads = e->cfg.heads, dh = H / heads;
  const int n_tok = batch * seq;
  const int M = (n_tok + 255) / 256 * 256;
  const int wpb = 8;
  __nv_bfloat16* x = e->x.as<__nv_bfloat16>();
  __nv_bfloat16* y = e->y.as<__nv_bfloat16>();
  // Deferred LayerNorm (hidden sizes without the fused GEMM + LayerNorm kernel; option defer_ln=0 disables): no
  // LayerNorm output is ever written -- see EPI_ROWFOLD in encoder_kernels.cuh.  The x buffer then holds the
  // un-normalised input / output of a layer, the y buffer the one between its two blocks, each with its statistic slots.
  const bool defer = e->defer_capable && opt(OPT_DEFER_LN) != 0 && opt(OPT_GEMM_PAIR) != 0;
  const int ln_parts = H / GEMM_OUT_CHUNK;
  float* stats_x = e->ln_stats0.as<float>();
  float* stats_y = e->ln_stats1.as<float>();
  embed_ln_kernel<NV><<<(n_tok + wpb - 1) / wpb, wpb * 32, 0, st>>>(
      d_ids, n_tok, seq, e->cfg.vocab_size, e->word.as<__nv_bfloat16>(), e->pos.as<__nv_bfloat16>(),
      e->type0.as<__nv_bfloat16>(), e->emb_g.as<float>(), e->emb_b.as<float>(), e->cfg.ln_eps, x);
  if (defer)   // layer 0 reads the normalised embedding rows through "identity" statistics
    ln_identity_stats_kernel<<<(unsigned)(((int64_t)M * ln_parts + 255) / 256), 256, 0, st>>>(stats_x, M, ln_parts, (float)H);
  B200_CUDA(cudaGetLastError());

  const size_t att_smem = 2 * (size_t)seq * (dh + 8) * 2;   // K and V tiles, rows padded by 16 bytes
  if (att_smem > 232448) B200_FAIL(B200_ERR_UNSUPPORTED, "sequence length %d needs %zu bytes of shared memory", seq, att_smem);
  if (dh == 32) B200_CUDA(cudaFuncSetAttribute(attention_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)att_smem));
  else B200_CUDA(cudaFuncSetAttribute(attention_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)att_smem));

  // 32-wide heads and sequences up to 256 tokens (MiniLM-class) run on the tcgen05 attention kernel;
  // other shapes use the mma.sync one.  Option attention_tc=0 forces the latter (A/B timing).
  const bool use_att_tc = dh == ATT_DH && seq <= ATT_MAX_S && seq % 32 == 0 && opt(OPT_ATTENTION_TC) != 0;
  const size_t att_tc_smem = ATT_SMEM_BYTES;   // alignment slack, two stages, barriers + exchange buffers
  const int att_exact_max = opt(OPT_ATT_EXACT_MAX);
  CUtensorMap map_att_q, map_att_kv, map_att_out;
  if (use_att_tc) {
    B200_TRY(make_tma_2d_bf16_plain(&map_att_out, e->ctx.p, (uint64_t)e->m_cap, (uint64_t)H, 32, ATT_DH / 2));   // one warp's rows x dims
    B200_TRY(make_tma_2d_bf16_sw64(&map_att_q, e->qkv.p, (uint64_t)e->m_cap, (uint64_t)(3 * H), (uint32_t)std::min(seq, 128), ATT_DH));
    B200_TRY(make_tma_2d_bf16_sw64(&map_att_kv, e->qkv.p, (uint64_t)e->m_cap, (uint64_t)(3 * H), (uint32_t)seq, ATT_DH));
  }
  // how many of every 8 exponentials the attention kernel computes on the FMA pipe (option att_poly)
  typedef void (*AttKernel)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const AttParams);
  static const AttKernel att_kernels[5] = {attention_tc_kernel<0>, attention_tc_kernel<1>, attention_tc_kernel<2>,
                                           attention_tc_kernel<3>, attention_tc_kernel<4>};
  const AttKernel att_kernel = att_kernels[std::min(std::max(opt(OPT_ATT_POLY), 0), 4)];
  if (use_att_tc) B200_CUDA(cudaFuncSetAttribute(att_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)att_tc_smem));

  // hidden 384 = two 192-column tiles: residual + LayerNorm fused into the producing GEMM (option fuse_ln=0 disables)
  const bool pair_ok = opt(OPT_GEMM_PAIR) != 0;
  const bool fuse_ffn = H == FFN_H && M % 256 == 0 && I % FFN_SLICE == 0 && opt(OPT_FUSE_FFN) != 0 && pair_ok;
  const bool fuse_ln = H == 384 && M % 256 == 0 && opt(OPT_FUSE_LN) != 0 && pair_ok;
  // serpentine row order: every kernel walks the rows in the opposite direction of the one before it
  // (option serpentine=0 disables)
  const bool serp = opt(OPT_SERPENTINE) != 0;
  // 64-wide heads over whole 128-key blocks (bert-base / mpnet-base-class shapes): the block-wise tcgen05 kernel
  const bool use_att_tc64 = dh == A64_DH && seq % A64_BK == 0 && opt(OPT_ATTENTION_TC) != 0;
  typedef void (*Att64Kernel)(const CUtensorMap, const CUtensorMap, const Att64Params);
  const Att64Kernel att64_kernel = opt(OPT_ATT_POLY) > 0 ? attention_tc64_kernel<1> : attention_tc64_kernel<0>;
  CUtensorMap map_a64_qkv, map_a64_out;
  if (use_att_tc64) {
    B200_TRY(make_tma_2d_bf16(&map_a64_qkv, e->qkv.p, (uint64_t)e->m_cap, (uint64_t)(3 * H), 128, A64_DH));
    B200_TRY(make_tma_2d_bf16_plain(&map_a64_out, e->ctx.p, (uint64_t)e->m_cap, (uint64_t)H, 32, 32));
    B200_CUDA(cudaFuncSetAttribute(att64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)A64_SMEM_BYTES));
  }
  int kcount = 0;
  for (int l = 0; l < e->cfg.layers; ++l) {
    LayerWeights* L = e->layers[l];
    GemmParams p{};
    // Q,K,V projections in one GEMM: qkv = x Wqkv^T + bqkv
    p.M = M; p.N = 3 * H; p.K = H; p.bias = L->bqkv.as<float>(); p.reverse = serp ? (kcount++ & 1) : 0;
    // 64-wide heads on whole CTA-pair tiles: the epilogue also folds the per-(sequence, head) maxima of |q|^2 and |k|^2
    // that attention_tc64 takes its softmax reference from (option att_norms_fused=0: a separate pass over Q and K)
    const bool fold_norms = use_att_tc64 && opt(OPT_ATT_NORMS_FUSED) != 0 && M % (2 * GEMM_BLOCK_M) == 0 &&
                            pick_block_n(3 * H) == 192 && opt(OPT_GEMM_PAIR) != 0;
    // the LayerNorm in front of this layer (deferred form): none for layer 0 (the embedding's has been applied), else the
    // previous layer's second one
    const float* in_g = l == 0 ? e->ln_ones.as<float>() : e->layers[l - 1]->ln2_g.as<float>();
    const float* in_b = l == 0 ? e->ln_zeros.as<float>() : e->layers[l - 1]->ln2_b.as<float>();
    if (fold_norms) {
      B200_CUDA(cudaMemsetAsync(e->att_kmax2.p, 0, (size_t)2 * batch * heads * 4, st));
      p.hn_k2 = e->att_kmax2.as<float>(); p.hn_q2 = p.hn_k2 + (size_t)batch * heads;
      p.hn_S = seq; p.hn_H = H; p.hn_heads = heads; p.hn_rows = n_tok;
    }
    if (defer) {
      p.bias = l == 0 ? L->bqkv.as<float>() : L->bqkv_g.as<float>(); p.fold_c = L->cqkv.as<float>(); p.eps = e->cfg.ln_eps;
      p.ln_stats_in = stats_x; p.ln_parts = ln_parts; p.ln_inv_n = 1.0f / (float)H;
      const GemmIO io{&e->map_x, &L->map_wqkv, l == 0 ? &L->map_wqkv_h : &L->map_wqkv_gh, &e->map_qkv_st, &e->map_qkv_st};
      if (fold_norms) B200_TRY((launch_gemm_t<192, EPI_ROWFOLD | EPI_HEADNORM, 2>(e, io, p, st)));
      else B200_TRY((launch_gemm_t<192, EPI_ROWFOLD, 2>(e, io, p, st)));
    } else if (fold_norms) {
      B200_TRY((launch_gemm_t<192, EPI_HEADNORM, 2>(e, GemmIO{&e->map_x, &L->map_wqkv, &L->map_wqkv_h, &e->map_qkv_st, &e->map_qkv_st}, p, st)));
    } else {
      B200_TRY(launch_gemm<EPI_BIAS>(e, GemmIO{&e->map_x, &L->map_wqkv, &L->map_wqkv_h, &e->map_qkv_st, &e->map_qkv_st}, p, st));
    }
    if (use_att_tc) {
      AttParams ap;
      ap.S = seq; ap.H = H; ap.heads = heads; ap.n_mt = (seq + 127) / 128; ap.n_units = batch * heads * ap.n_mt;
      ap.lens = d_lens; ap.exact_max = att_exact_max; ap.reverse = serp ? (kcount++ & 1) : 0;
      ap.no_turns = opt(OPT_ATT_TURNS) == 0;
      B200_PROF("attention", 4.0 * batch * (double)seq * seq * H, 2.0 * 4.0 * n_tok * H, st);
      cudaLaunchConfig_t cfg{};
      cfg.gridDim = dim3(std::min(ap.n_units, e->num_sms));
      cfg.blockDim = dim3(ATT_THREADS);
      cfg.dynamicSmemBytes = att_tc_smem;
      cfg.stream = st;
      cudaLaunchAttribute attr[2];
      cfg.attrs = attr; cfg.numAttrs = hot_launch_attrs(attr, false);
      B200_CUDA(cudaLaunchKernelEx(&cfg, att_kernel, map_att_q, map_att_kv, map_att_out, ap));
    } else if (use_att_tc64) {
      B200_CUDA(cudaMemsetAsync(e->att_flag.p, 0, 4, st));
      if (fold_norms)
        att_flag_kernel<<<(batch * heads + 255) / 256, 256, 0, st>>>(e->att_kmax2.as<float>() + (size_t)batch * heads,
                                                                    e->att_kmax2.as<float>(), batch * heads, e->att_flag.as<int32_t>());
      else
        att_norms_kernel<A64_DH><<<dim3(heads, batch), 256, 0, st>>>(e->qkv.as<__nv_bfloat16>(), seq, H, heads,
                                                                    e->att_kmax2.as<float>(), e->att_flag.as<int32_t>());
      B200_CUDA(cudaGetLastError());
      Att64Params ap{};
      ap.S = seq; ap.H = H; ap.heads = heads; ap.n_mt = seq / 128; ap.n_units = batch * heads * ap.n_mt;
      ap.lens = d_lens; ap.kmax2 = e->att_kmax2.as<float>(); ap.exact_flag = e->att_flag.as<int32_t>();
      ap.force_exact = att_exact_max; ap.reverse = serp ? (kcount++ & 1) : 0;
      B200_PROF("attention", 4.0 * batch * (double)seq * seq * H, 2.0 * 4.0 * n_tok * H, st);
      cudaLaunchConfig_t cfg{};
      cfg.gridDim = dim3(std::min(ap.n_units, e->num_sms));
      cfg.blockDim = dim3(A64_THREADS);
      cfg.dynamicSmemBytes = A64_SMEM_BYTES;
      cfg.stream = st;
      cudaLaunchAttribute attr[2];
      cfg.attrs = attr; cfg.numAttrs = 0;
      B200_CUDA(cudaLaunchKernelEx(&cfg, att64_kernel, map_a64_qkv, map_a64_out, ap));
    } else {
      B200_PROF("attention", 4.0 * batch * (double)seq * seq * H, 2.0 * 4.0 * n_tok * H, st);
      if (dh == 32)
        attention_kernel<32><<<dim3(heads, batch), 256, att_smem, st>>>(e->qkv.as<__nv_bfloat16>(), d_lens, seq, H,
                                                                        e->ctx.as<__nv_bfloat16>());
      else
        attention_kernel<64><<<dim3(heads, batch), 256, att_smem, st>>>(e->qkv.as<__nv_bfloat16>(), d_lens, seq, H,
                                                                        e->ctx.as<__nv_bfloat16>());
    }
    B200_CUDA(cudaGetLastError());
    // y = ctx Wo^T + bo + x ; x = LN(y)
    p.N = H; p.K = H; p.bias = L->bo.as<float>(); p.reverse = serp ? (kcount++ & 1) : 0;
    if (defer) {
      // y = ctx Wo^T + bo + LN_in(x buffer): the residual is rebuilt from the un-normalised rows; statistics of y go out
      p.gamma = in_g; p.beta = in_b; p.eps = e->cfg.ln_eps;
      p.ln_stats_in = stats_x; p.ln_parts = ln_parts; p.ln_inv_n = 1.0f / (float)H; p.ln_stats_out = stats_y;
      B200_TRY((launch_gemm_t<192, EPI_BIAS_RESID | EPI_RESID_LN | EPI_STATS, 2>(
          e, GemmIO{&e->map_ctx, &L->map_wo, &L->map_wo_h, &e->map_y_st, &e->map_x_res}, p, st)));
    } else if (fuse_ln) {
      p.gamma = L->ln1_g.as<float>(); p.beta = L->ln1_b.as<float>(); p.eps = e->cfg.ln_eps;
      B200_TRY(launch_gemm_ln(e, GemmIO{&e->map_ctx, &L->map_wo, &L->map_wo_h, &e->map_x_res, &e->map_x_res}, p, st));
    } else {
      B200_TRY(launch_gemm<EPI_BIAS_RESID>(e, GemmIO{&e->map_ctx, &L->map_wo, &L->map_wo_h, &e->map_y_st, &e->map_x_res}, p, st));
      layernorm_kernel<NV><<<(n_tok + wpb - 1) / wpb, wpb * 32, 0, st>>>(y, n_tok, L->ln1_g.as<float>(), L->ln1_b.as<float>(),
                                                                        e->cfg.ln_eps, x);
      B200_CUDA(cudaGetLastError());
    }
    // ffn = gelu(x W1^T + b1) ; y = ffn W2^T + b2 + x ; x = LN(y)
    if (fuse_ffn) {
      B200_TRY(launch_ffn_fused(e, L, M, I, serp ? (kcount++ & 1) : 0, st));
      continue;
    }
    if (defer) {
      // ffn = gelu(LN1(y) W1^T + b1) from the un-normalised y; then x buffer = ffn W2^T + b2 + LN1(y), with its statistics
      p = GemmParams{};
      p.M = M; p.N = I; p.K = H; p.bias = L->b1_g.as<float>(); p.fold_c = L->c1.as<float>(); p.eps = e->cfg.ln_eps;
      p.ln_stats_in = stats_y; p.ln_parts = ln_parts; p.ln_inv_n = 1.0f / (float)H; p.reverse = serp ? (kcount++ & 1) : 0;
      B200_TRY((launch_gemm_t<192, EPI_ROWFOLD | EPI_BIAS_GELU, 2>(
          e, GemmIO{&e->map_y, &L->map_w1, &L->map_w1_gh, &e->map_ffn_st, &e->map_ffn_st}, p, st)));
      p = GemmParams{};
      p.M = M; p.N = H; p.K = I; p.bias = L->b2.as<float>(); p.reverse = serp ? (kcount++ & 1) : 0;
      p.gamma = L->ln1_g.as<float>(); p.beta = L->ln1_b.as<float>(); p.eps = e->cfg.ln_eps;
      p.ln_stats_in = stats_y; p.ln_parts = ln_parts; p.ln_inv_n = 1.0f / (float)H; p.ln_stats_out = stats_x;
      B200_TRY((launch_gemm_t<192, EPI_BIAS_RESID | EPI_RESID_LN | EPI_STATS, 2>(
          e, GemmIO{&e->map_ffn, &L->map_w2, &L->map_w2_h, &e->map_x_res, &e->map_y_st}, p, st)));
      continue;
    }
    p.N = I; p.K = H; p.bias = L->b1.as<float>(); p.reverse = serp ? (kcount++ & 1) : 0;
    B200_TRY(launch_gemm<EPI_BIAS_GELU>(e, GemmIO{&e->map_x, &L->map_w1, &L->map_w1_h, &e->map_ffn_st, &e->map_ffn_st}, p, st));
    p.N = H; p.K = I; p.bias = L->b2.as<float>(); p.reverse = serp ? (kcount++ & 1) : 0;
    if (fuse_ln) {
      p.gamma = L->ln2_g.as<float>(); p.beta = L->ln2_b.as<float>(); p.eps = e->cfg.ln_eps;
      B200_TRY(launch_gemm_ln(e, GemmIO{&e->map_ffn, &L->map_w2, &L->map_w2_h, &e->map_x_res, &e->map_x_res}, p, st));
    } else {
      B200_TRY(launch_gemm<EPI_BIAS_RESID>(e, GemmIO{&e->map_ffn, &L->map_w2, &L->map_w2_h, &e->map_y_st, &e->map_x_res}, p, st));
      layernorm_kernel<NV><<<(n_tok + wpb - 1) / wpb, wpb * 32, 0, st>>>(y, n_tok, L->ln2_g.as<float>(), L->ln2_b.as<float>(),
                                                                        e->cfg.ln_eps, x);
      B200_CUDA(cudaGetLastError());
    }
  }
  if (defer) {   // the last layer's second LayerNorm, in place (a row is read whole before it is written)
    const LayerWeights* L = e->layers.back();
    layernorm_kernel<NV><<<(n_tok + wpb - 1) / wpb, wpb * 32, 0, st>>>(x, n_tok, L->ln2_g.as<float>(), L->ln2_b.as<float>(),
                                                                      e->cfg.ln_eps, x);
    B200_CUDA(cudaGetLastError());
  }
  pool_normalize_kernel<<<batch, 256, 0, st>>>(x, d_lens, seq, H, d_out);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

extern "C" {

int b200_encoder_create(const b200_encoder_config_t* cfg, int device, b200_encoder_t** out) {
  if (!out) B200_FAIL(B200_ERR_INVALID, "out is NULL");
  *out = nullptr;
  if (!cfg) B200_FAIL(B200_ERR_INVALID, "config is NULL");
  if (cfg->hidden <= 0 || cfg->hidden % 128 != 0 || cfg->hidden > 1024)
    B200_FAIL(B200_ERR_UNSUPPORTED, "hidden=%d: must be a multiple of 128 up to 1024", cfg->hidden);
  if (cfg->heads <= 0 || cfg->hidden % cfg->heads != 0)
    B200_FAIL(B200_ERR_INVALID, "hidden=%d is not divisible by heads=%d", cfg->hidden, cfg->heads);
  const int dh = cfg->hidden / cfg->heads;
  if (dh != 32 && dh != 64) B200_FAIL(B200_ERR_UNSUPPORTED, "head dimension %d: only 32 and 64 are implemented", dh);
  if (cfg->intermediate <= 0 || cfg->intermediate % 64 != 0)
    B200_FAIL(B200_ERR_UNSUPPORTED, "intermediate=%d must be a multiple of 64", cfg->intermediate);
  if (cfg->layers <= 0 || cfg->vocab_size <= 0 || cfg->max_pos <= 0 || cfg->type_vocab <= 0)
    B200_FAIL(B200_ERR_INVALID, "layers, vocab_size, max_pos and type_vocab must be positive");
  B200_TRY(check_device_is_sm100(device));
  b200_encoder* e = new b200_encoder();
  e->cfg = *cfg;
  e->device = device;
  cudaDeviceGetAttribute(&e->num_sms, cudaDevAttrMultiProcessorCount, device);
  {
    DeviceGuard g(device);
    const int rc = e->order.init();
    if (rc != B200_OK) {
      delete e;
      return rc;
    }
  }
  *out = e;
  return B200_OK;
}

int b200_encoder_destroy(b200_encoder_t* e) {
  if (!e) return B200_OK;
  DeviceGuard g(e->device);
  cudaDeviceSynchronize();
  delete e;
  return B200_OK;
}

int b200_encoder_hidden(const b200_encoder_t* e) { return e ? e->cfg.hidden : 0; }

int b200_encoder_load_tensor(b200_encoder_t* e, const char* name, const float* data, int64_t numel) {
  if (!e || !name || !data || numel <= 0) B200_FAIL(B200_ERR_INVALID, "bad load_tensor arguments");
  std::lock_guard<std::mutex> lk(e->mu);
  if (e->finalized) B200_FAIL(B200_ERR_INVALID, "encoder is already finalized");
  e->pending[name].assign(data, data + numel);
  return B200_OK;
}

int b200_encoder_finalize(b200_encoder_t* e) {
  if (!e) B200_FAIL(B200_ERR_INVALID, "encoder is NULL");
  std::lock_guard<std::mutex> lk(e->mu);
  if (e->finalized) return B200_OK;
  DeviceGuard g(e->device);
  if (!g.ok) B200_FAIL(B200_ERR_CUDA, "cudaSetDevice(%d) failed", e->device);
  const size_t H = e->cfg.hidden, I = e->cfg.intermediate;
  const std::vector<float>* t = nullptr;
  B200_TRY(take(e, "embeddings.word_embeddings.weight", (size_t)e->cfg.vocab_size * H, &t));
  B200_TRY(upload_bf16(e->word, t->data(), t->size()));
  B200_TRY(take(e, "embeddings.position_embeddings.weight", (size_t)e->cfg.max_pos * H, &t));
  B200_TRY(upload_bf16(e->pos, t->data(), t->size()));
  B200_TRY(take(e, "embeddings.token_type_embeddings.weight", (size_t)e->cfg.type_vocab * H, &t));
  B200_TRY(upload_bf16(e->type0, t->data(), H));  // token type ids are all zero on this path
  B200_TRY(take(e, "embeddings.LayerNorm.weight", H, &t));
  B200_TRY(upload_f32(e->emb_g, t->data(), H));
  B200_TRY(take(e, "embeddings.LayerNorm.bias", H, &t));
  B200_TRY(upload_f32(e->emb_b, t->data(), H));
  // deferred LayerNorm needs whole 192-column tiles everywhere and an even number of 64-column statistic slots;
  // hidden 384 has the fused GEMM + LayerNorm kernel instead
  e->defer_capable = H != 384 && H % 192 == 0 && (H / 64) % 2 == 0 && I % 192 == 0;
  // the LayerNorm in front of the layer being loaded; none for layer 0 (the embedding's is applied by embed_ln_kernel)
  std::vector<float> prev_g(H, 1.0f), prev_b(H, 0.0f);
  if (e->defer_capable) {
    B200_TRY(upload_f32(e->ln_ones, prev_g.data(), H));
    B200_TRY(upload_f32(e->ln_zeros, prev_b.data(), H));
  }
  // W (.) gamma as bf16, its row sums, b + W beta
  auto fold = [&](const float* w, const float* b, size_t N, const std::vector<float>& g, const std::vector<float>& be,
                  DevBuf& wg, DevBuf& c, DevBuf& bg) -> int {
    std::vector<uint16_t> wq(N * H);
    std::vector<float> cs(N), bf(N);
    for (size_t n = 0; n < N; ++n) {
      double csum = 0.0, bsum = 0.0;
      for (size_t k = 0; k < H; ++k) {
        const uint16_t q = f32_to_bf16_host(w[n * H + k] * g[k]);
        wq[n * H + k] = q;
        const uint32_t u = (uint32_t)q << 16;
        float back;
        memcpy(&back, &u, 4);
        csum += back;
        bsum += (double)w[n * H + k] * be[k];
      }
      cs[n] = (float)csum;
      bf[n] = b[n] + (float)bsum;
    }
    B200_TRY(wg.alloc(N * H * 2));
    B200_CUDA(cudaMemcpy(wg.p, wq.data(), N * H * 2, cudaMemcpyHostToDevice));
    B200_TRY(upload_f32(c, cs.data(), N));
    B200_TRY(upload_f32(bg, bf.data(), N));
    return B200_OK;
  };
  for (int l = 0; l < e->cfg.layers; ++l) {
    const std::string p = "encoder.layer." + std::to_string(l) + ".";
    LayerWeights* L = new LayerWeights();
    e->layers.push_back(L);
    std::vector<float> wqkv(3 * H * H), bqkv(3 * H);
    const char* names[3] = {"query", "key", "value"};
    for (int j = 0; j < 3; ++j) {
      B200_TRY(take(e, p + "attention.self." + names[j] + ".weight", H * H, &t));
      memcpy(wqkv.data() + (size_t)j * H * H, t->data(), H * H * 4);
      B200_TRY(take(e, p + "attention.self." + names[j] + ".bias", H, &t));
      memcpy(bqkv.data() + (size_t)j * H, t->data(), H * 4);
    }
    // The attention kernels take 2^x of their scores: fold log2(e) / sqrt(head dim) into the query
    // projection here, in fp32 before the bf16 rounding, so that Q K^T needs no per-score multiply.
    {
      const float c = 1.4426950408889634f / sqrtf((float)(H / e->cfg.heads));
      for (size_t i = 0; i < (size_t)H * H; ++i) wqkv[i] *= c;
      for (size_t i = 0; i < (size_t)H; ++i) bqkv[i] *= c;
    }
    B200_TRY(upload_bf16(L->wqkv, wqkv.data(), wqkv.size()));
    B200_TRY(upload_f32(L->bqkv, bqkv.data(), bqkv.size()));
    if (e->defer_capable) {
      B200_TRY(fold(wqkv.data(), bqkv.data(), 3 * H, prev_g, prev_b, L->wqkv_g, L->cqkv, L->bqkv_g));
      B200_TRY(weight_map(&L->map_wqkv_gh, L->wqkv_g, (int)(3 * H), (int)H, 2));
    }
    B200_TRY(take(e, p + "attention.output.dense.weight", H * H, &t));
    B200_TRY(upload_bf16(L->wo, t->data(), t->size()));
    B200_TRY(take(e, p + "attention.output.dense.bias", H, &t));
    B200_TRY(upload_f32(L->bo, t->data(), H));
    B200_TRY(take(e, p + "attention.output.LayerNorm.weight", H, &t));
    B200_TRY(upload_f32(L->ln1_g, t->data(), H));
    B200_TRY(take(e, p + "attention.output.LayerNorm.bias", H, &t));
    B200_TRY(upload_f32(L->ln1_b, t->data(), H));
    B200_TRY(take(e, p + "intermediate.dense.weight", I * H, &t));
    B200_TRY(upload_bf16(L->w1, t->data(), t->size()));
    const std::vector<float>* tb1 = nullptr;
    B200_TRY(take(e, p + "intermediate.dense.bias", I, &tb1));
    B200_TRY(upload_f32(L->b1, tb1->data(), I));
    if (e->defer_capable) {
      const std::vector<float>*g1 = nullptr, *be1 = nullptr;
      B200_TRY(take(e, p + "attention.output.LayerNorm.weight", H, &g1));
      B200_TRY(take(e, p + "attention.output.LayerNorm.bias", H, &be1));
      B200_TRY(fold(t->data(), tb1->data(), I, *g1, *be1, L->w1_g, L->c1, L->b1_g));
      B200_TRY(weight_map(&L->map_w1_gh, L->w1_g, (int)I, (int)H, 2));
    }
    B200_TRY(take(e, p + "output.dense.weight", H * I, &t));
    B200_TRY(upload_bf16(L->w2, t->data(), t->size()));
    B200_TRY(take(e, p + "output.dense.bias", H, &t));
    B200_TRY(upload_f32(L->b2, t->data(), H));
    B200_TRY(take(e, p + "output.LayerNorm.weight", H, &t));
    B200_TRY(upload_f32(L->ln2_g, t->data(), H));
    prev_g = *t;
    B200_TRY(take(e, p + "output.LayerNorm.bias", H, &t));
    B200_TRY(upload_f32(L->ln2_b, t->data(), H));
    prev_b = *t;
    B200_TRY(weight_map(&L->map_wqkv, L->wqkv, (int)(3 * H), (int)H));
    B200_TRY(weight_map(&L->map_wo, L->wo, (int)H, (int)H));
    B200_TRY(weight_map(&L->map_w1, L->w1, (int)I, (int)H));
    B200_TRY(weight_map(&L->map_w2, L->w2, (int)H, (int)I));
    B200_TRY(weight_map(&L->map_wqkv_h, L->wqkv, (int)(3 * H), (int)H, 2));
    B200_TRY(weight_map(&L->map_wo_h, L->wo, (int)H, (int)H, 2));
    B200_TRY(weight_map(&L->map_w1_h, L->w1, (int)I, (int)H, 2));
    B200_TRY(weight_map(&L->map_w2_h, L->w2, (int)H, (int)I, 2));
    B200_TRY(make_tma_2d_bf16(&L->map_w1_f, L->w1.p, (uint64_t)I, (uint64_t)H, FFN_SLICE / 2, GEMM_BLOCK_K));
  }
  e->pending.clear();
  e->finalized = true;
  return B200_OK;
}

int b200_encode(b200_encoder_t* e, const int32_t* input_ids, const int32_t* lengths, int batch, int seq,
                int inputs_on_device, float* out, int out_on_device, void* stream) {
  if (!e) B200_FAIL(B200_ERR_INVALID, "encoder is NULL");
  if (batch < 0 || seq < 0) B200_FAIL(B200_ERR_INVALID, "negative batch / seq");
  if (batch == 0) return B200_OK;
  if (!input_ids || !lengths || !out) B200_FAIL(B200_ERR_INVALID, "NULL buffer");
  if (seq == 0 || seq % 16 != 0) B200_FAIL(B200_ERR_INVALID, "seq=%d must be a positive multiple of 16 (pad with id 0)", seq);
  if (seq > e->cfg.max_pos) B200_FAIL(B200_ERR_INVALID, "seq=%d exceeds max_position_embeddings=%d", seq, e->cfg.max_pos);
  std::lock_guard<std::mutex> lk(e->mu);
  if (!e->finalized) B200_FAIL(B200_ERR_INVALID, "encoder weights are not finalized");
  DeviceGuard g(e->device);
  if (!g.ok) B200_FAIL(B200_ERR_CUDA, "cudaSetDevice(%d) failed", e->device);
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t n_tok = (int64_t)batch * seq;
  if (n_tok > (int64_t)1 << 30) B200_FAIL(B200_ERR_UNSUPPORTED, "batch*seq too large");
  B200_TRY(e->order.enter(st));
  B200_TRY(ensure_activations(e, n_tok, batch, st));
  const int32_t* d_ids = input_ids;
  const int32_t* d_lens = lengths;
  if (!inputs_on_device) {
    B200_CUDA(cudaMemcpyAsync(e->ids.p, input_ids, (size_t)n_tok * 4, cudaMemcpyHostToDevice, st));
    B200_CUDA(cudaMemcpyAsync(e->lens.p, lengths, (size_t)batch * 4, cudaMemcpyHostToDevice, st));
    d_ids = e->ids.as<int32_t>();
    d_lens = e->lens.as<int32_t>();
  }
  float* d_out = out_on_device ? out : e->out.as<float>();
  int rc;
  switch (e->cfg.hidden / 128) {
    case 1: rc = run_encoder<1>(e, d_ids, d_lens, batch, seq, d_out, st); break;
    case 2: rc = run_encoder<2>(e, d_ids, d_lens, batch, seq, d_out, st); break;
    case 3: rc = run_encoder<3>(e, d_ids, d_lens, batch, seq, d_out, st); break;
    case 4: rc = run_encoder<4>(e, d_ids, d_lens, batch, seq, d_out, st); break;
    case 6: rc = run_encoder<6>(e, d_ids, d_lens, batch, seq, d_out, st); break;
    case 8: rc = run_encoder<8>(e, d_ids, d_lens, batch, seq, d_out, st); break;
    default: B200_FAIL(B200_ERR_UNSUPPORTED, "hidden=%d is not instantiated (128,256,384,512,768,1024)", e->cfg.hidden);
  }
  B200_TRY(rc);
  if (!out_on_device) B200_CUDA(cudaMemcpyAsync(out, d_out, (size_t)batch * e->cfg.hidden * 4, cudaMemcpyDeviceToHost, st));
  B200_TRY(e->order.leave(st));
  if (!out_on_device) B200_CUDA(cudaStreamSynchronize(st));
  return B200_OK;
}

int b200_encoder_peek(b200_encoder_t* e, int which, float* out_host, int64_t numel) {
  if (!e || !out_host || numel < 0) B200_FAIL(B200_ERR_INVALID, "bad peek arguments");
  std::lock_guard<std::mutex> lk(e->mu);
  DeviceGuard g(e->device);
  DevBuf* bufs[5] = {&e->x, &e->qkv, &e->ctx, &e->y, &e->ffn};
  if (which < 0 || which > 4) B200_FAIL(B200_ERR_INVALID, "which must be 0..4");
  if ((size_t)numel * 2 > bufs[which]->bytes) B200_FAIL(B200_ERR_INVALID, "numel exceeds the buffer");
  std::vector<uint16_t> tmp((size_t)numel);
  B200_CUDA(cudaDeviceSynchronize());
  B200_CUDA(cudaMemcpy(tmp.data(), bufs[which]->p, (size_t)numel * 2, cudaMemcpyDeviceToHost));
  for (int64_t i = 0; i < numel; ++i) {
    const uint32_t u = (uint32_t)tmp[(size_t)i] << 16;
    memcpy(&out_host[i], &u, 4);
  }
  return B200_OK;
}

}  // extern "C"

#ifdef B200_ATT_TIMING
extern "C" int b200_debug_att_timing(unsigned long long* out) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, b200::g_att_timing, 16 * sizeof(unsigned long long));
  unsigned long long z[16] = {0};
  cudaMemcpyToSymbol(b200::g_att_timing, z, sizeof(z));
  return 0;
}
#endif

#ifdef B200_GEMM_TIMING
extern "C" int b200_debug_gemm_timing(unsigned long long* out) {
  const unsigned long long flag = out[7];   // the caller passes the experiment flag in through out[7]
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, b200::g_gemm_timing, 8 * sizeof(unsigned long long));
  unsigned long long z[8] = {0};
  z[7] = flag;
  cudaMemcpyToSymbol(b200::g_gemm_timing, z, sizeof(z));
  return 0;
}
#endif
