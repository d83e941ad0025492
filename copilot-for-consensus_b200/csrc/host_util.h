// host_util.h -- host-side plumbing shared by index.cu and encoder.cu: thread-local error
// reporting for the C ABI, CUDA error checks, device buffers, TMA descriptor encoding.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>

#include "../../include/b200_retrieval.h"

namespace b200 {

void set_error(const char* fmt, ...);
const char* get_error();

#define B200_FAIL(code, ...)        \
  do {                              \
    ::b200::set_error(__VA_ARGS__); \
    return (code);                  \
  } while (0)

#define B200_CUDA(expr)                                                                          \
  do {                                                                                           \
    cudaError_t _e = (expr);                                                                     \
    if (_e != cudaSuccess) {                                                                     \
      ::b200::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return _e == cudaErrorMemoryAllocation ? B200_ERR_NOMEM : B200_ERR_CUDA;                   \
    }                                                                                            \
  } while (0)

#define B200_TRY(expr)          \
  do {                          \
    int _rc = (expr);           \
    if (_rc != B200_OK) return _rc; \
  } while (0)

// Makes `device` current for the scope of a call and restores the caller's device afterwards.
struct DeviceGuard {
  int prev = -1;
  bool ok = false;
  explicit DeviceGuard(int device) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    ok = cudaSetDevice(device) == cudaSuccess;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

// Owning device allocation that can grow (contents preserved up to `keep` bytes).
struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  int alloc(size_t n) {
    release();
    if (n == 0) return B200_OK;
    B200_CUDA(cudaMalloc(&p, n));
    bytes = n;
    return B200_OK;
  }
  int grow(size_t n, size_t keep, cudaStream_t st) {
    if (n <= bytes) return B200_OK;
    void* np = nullptr;
    B200_CUDA(cudaMalloc(&np, n));
    if (p && keep) {
      cudaError_t e = cudaMemcpyAsync(np, p, keep, cudaMemcpyDeviceToDevice, st);
      if (e == cudaSuccess) e = cudaStreamSynchronize(st);
      if (e != cudaSuccess) {
        cudaFree(np);
        B200_FAIL(B200_ERR_CUDA, "device buffer growth copy failed: %s", cudaGetErrorString(e));
      }
    }
    if (p) cudaFree(p);
    p = np;
    bytes = n;
    return B200_OK;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    bytes = 0;
  }
  ~DevBuf() { release(); }
  DevBuf() = default;
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  template <typename T>
  T* as() const { return reinterpret_cast<T*>(p); }
};

// Orders the GPU work of successive calls on one handle when they arrive on different streams.  Every
// handle owns one set of scratch buffers; the handle's mutex serialises the host side of the calls, but
// with device buffers a call returns while its kernels are still queued, so a second call on another
// stream could otherwise run over the first one's scratch.  enter() makes the new stream wait for the
// previous call's last kernel (a no-op when the stream is the same), leave() marks the end of this call.
struct StreamOrder {
  cudaEvent_t ev = nullptr;
  cudaStream_t last = nullptr;
  bool used = false;
  int init() {
    B200_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    return B200_OK;
  }
  int enter(cudaStream_t st) {
    if (used && last != st) B200_CUDA(cudaStreamWaitEvent(st, ev, 0));
    return B200_OK;
  }
  int leave(cudaStream_t st) {
    B200_CUDA(cudaEventRecord(ev, st));
    last = st;
    used = true;
    return B200_OK;
  }
  ~StreamOrder() {
    if (ev) cudaEventDestroy(ev);
  }
};

// 2-D 16-bit row-major tensor [rows, cols] -> TMA descriptor with a {box_cols, box_rows} box and
// 128-byte swizzle (box_cols must be 64 bf16 = 128 bytes).  Out-of-bounds elements read as zero.
int make_tma_2d_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows,
                     uint32_t box_cols);

// Same for a 64-byte swizzle (box_cols = 32 bf16), used by the attention kernel's Q/K/V tiles.
int make_tma_2d_bf16_sw64(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows,
                          uint32_t box_cols);

// Unswizzled box (rows of box_cols bf16, a multiple of 16 bytes), e.g. a warp's slice of an output tile.
int make_tma_2d_bf16_plain(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows,
                           uint32_t box_cols);

int check_device_is_sm100(int device);

// Process-wide switches between kernel variants (b200_set_option / b200_get_option in the C ABI).
// They exist for A/B timing and for the tests that drive every alternative path against the oracle;
// the library never reads the environment (the reference's adapters are pinned not to:
// adapters/copilot_vectorstore/tests/test_factory.py:75-107, scripts/check_no_runtime_env_vars.py:6-16).
enum Option {
  OPT_PDL = 0,            // 1: programmatic dependent launch between the encoder's kernels          (default 0)
  OPT_GEMM_PAIR,          // 0: single-CTA GEMM tiles instead of CTA pairs                            (default 1)
  OPT_GEMM_AST,           // 0: generic GEMM instead of the A-stationary one for K <= 384            (default 1)
  OPT_ATTENTION_TC,       // 0: mma.sync attention instead of the tcgen05 one                         (default 1)
  OPT_ATT_EXACT_MAX,      // 1: exact row-maximum pass for every softmax row                          (default 0)
  OPT_ATT_TURNS,          // 0: softmax groups free-running instead of taking turns on the SFU        (default 1)
  OPT_ATT_POLY,           // of every 8 exponentials, how many run on the FMA pipe instead of the SFU (0..4)  (default 1)
  OPT_FUSE_LN,            // 0: separate LayerNorm kernel                                             (default 1)
  OPT_FUSE_FFN,           // 0: FFN1 / FFN2 as two GEMMs                                              (default 1)
  OPT_SERPENTINE,         // 0: every kernel walks the rows in the same direction                     (default 1)
  OPT_SIM_PAIR,           // 0: single-CTA similarity tiles even with two or more query tiles        (default 1)
  OPT_SIM_BOUND,          // bit 0: cross-CTA bucket bound, bit 1: single-CTA stand-in bound          (default 3)
  OPT_SIM_REFRESH_MASK,   // the shared bounds are re-read every (mask + 1) tiles                     (default 0)
  OPT_SEARCH_CERTIFY,     // 0: skip the candidate-sufficiency certificate and its exhaustive re-run  (default 1)
  OPT_SIM_PAIR_KP,        // 16 / 20: candidates per query kept by the CTA-pair similarity kernel; 0 = by shard size (default 0)
  OPT_SIM_NSPLIT,         // > 0: force the number of corpus slices of the similarity kernel (A/B timing)             (default 0)
  OPT_SEARCH_SCREEN,      // 0: queries the certificate turns down go straight to the full exhaustive pass (no screened pass) (default 1)
  OPT_ATT_NORMS_FUSED,    // 0: a separate kernel re-reads Q and K for the softmax reference of 64-wide heads instead of the QKV epilogue (default 1)
  OPT_DEFER_LN,           // 0: hidden sizes without the fused GEMM + LayerNorm kernel run a separate LayerNorm pass instead of the deferred form (default 1)
  OPT_SIM_DRAIN,          // drain groups of the CTA-pair similarity kernel: 1 / 2 force, 0 = by shard size (default 0)
  OPT_COUNT
};
int opt(Option o);

// Optional per-kernel timing (off by default): when enabled through b200_profile_enable, each
// launch site brackets its kernel with a pair of CUDA events on the launching stream;
// b200_profile_read aggregates elapsed time, launch count and the caller-stated algorithmic
// flops / bytes per kernel name.  Used by bench.py for the roofline figures.
bool prof_enabled();
struct ProfScope {
  int slot = -1;
  cudaStream_t st;
  ProfScope(const char* name, double flops, double bytes, cudaStream_t st);
  ~ProfScope();
};
#define B200_PROF(name, flops, bytes, st) ::b200::ProfScope _prof_scope_##__LINE__(name, flops, bytes, st)

}  // namespace b200
