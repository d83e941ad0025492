// Micro-benchmark: cost per 32-element "chunk" of the softmax inner loop on one SM sub-partition.
//   mode 0: MUFU.EX2 only          mode 1: FFMA + MUFU.EX2 + FADD
//   mode 2: mode 1 + F2FP.BF16 pack   mode 3: mode 2 but half of the exponentials by polynomial (FMA pipe)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o sfu sfu.cu ; run: ./sfu
#include <cstdio>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../copilot-for-consensus_b200/csrc/common.cuh"
using namespace b200;

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// 2^x for x <= 0 on the FMA/ALU pipes: split x = n + f (f in [0,1)), degree-3 polynomial for 2^f,
// exponent added with an integer shift-add.
__device__ __forceinline__ float ex2_poly(float x) {
  x = fmaxf(x, -126.f);
  const float t = x + 12582912.f;                 // 1.5 * 2^23: round to nearest integer in the low bits
  const float n = t - 12582912.f;
  const float f = x - n;                          // [-0.5, 0.5]
  float pl = fmaf(f, 0.0555041f, 0.2402265f);
  pl = fmaf(pl, f, 0.6931472f);
  pl = fmaf(pl, f, 1.0f);
  return __int_as_float(__float_as_int(pl) + (__float_as_int(t) << 23));
}

template <int MODE>
__global__ void k(float* out, const float* in, int iters, long long* clk) {
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = in[(threadIdx.x * 32 + j) & 1023];
  float s0 = 0.f, s1 = 0.f;
  uint32_t acc = 0;
  const float scale = in[1024];
  float ms = in[1025];
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; ++it) {
    ms += 1e-4f;   // keeps the scaled arguments loop-variant
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      float p0, p1;
      if (MODE == 0) {
        p0 = ex2(v[2 * j]);
        p1 = ex2(v[2 * j + 1]);
        v[2 * j] = p0 - 1.5f;   // keeps a dependency without extra pipes mattering much
        v[2 * j + 1] = p1 - 1.5f;
      } else {
        const float a0 = fmaf(v[2 * j], scale, -ms), a1 = fmaf(v[2 * j + 1], scale, -ms);
        p0 = ex2(a0);
        p1 = (MODE == 3) ? ex2_poly(a1) : ex2(a1);
        s0 += p0;
        s1 += p1;
        if (MODE >= 2) {
          __nv_bfloat162 t = __floats2bfloat162_rn(p0, p1);
          acc ^= *reinterpret_cast<uint32_t*>(&t);
        }
      }
    }
  }
  const long long t1 = clock64();
  float r = s0 + s1 + __uint_as_float(acc);
#pragma unroll
  for (int j = 0; j < 32; ++j) r += v[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

// mode 4: mode 2 + tcgen05.st of the 16 packed words; mode 5: + tcgen05.ld of the 32 inputs (prefetched one chunk ahead)
template <int MODE>
__global__ void kt(float* out, const float* in, int iters, long long* clk) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(&slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = slot + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 64;
  uint32_t v[2][32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[0][j] = v[1][j] = __float_as_uint(in[(threadIdx.x * 32 + j) & 1023]);
  float s0 = 0.f, s1 = 0.f;
  const float scale = in[1024];
  float ms = in[1025];
  // seed TMEM with finite values
  {
    uint32_t z[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) z[j] = v[0][j];
    tmem_st16(tb, z); tmem_st16(tb + 16, z); tmem_st16(tb + 32, z); tmem_st16(tb + 48, z);
    tmem_st_wait();
  }
  __syncthreads();
  const long long t0 = clock64();
  if (MODE == 5) tmem_ld32(tb, v[0]);
#pragma unroll 1
  for (int it = 0; it < iters; it += 2) {
#pragma unroll
    for (int uu = 0; uu < 2; ++uu) {
      ms += 1e-4f;
      if (MODE == 5) {
        tmem_ld_wait_dep(v[uu]);
        tmem_ld32(tb, v[uu ^ 1]);
      }
      uint32_t pk[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const float a0 = fmaf(__uint_as_float(v[uu][2 * j]), scale, -ms), a1 = fmaf(__uint_as_float(v[uu][2 * j + 1]), scale, -ms);
        const float p0 = ex2(a0), p1 = ex2(a1);
        s0 += p0;
        s1 += p1;
        __nv_bfloat162 t = __floats2bfloat162_rn(p0, p1);
        pk[j] = *reinterpret_cast<uint32_t*>(&t);
      }
      tmem_st16(tb + 32, pk);
    }
  }
  tmem_st_wait();
  tmem_ld_wait();
  const long long t1 = clock64();
  float r = s0 + s1;
#pragma unroll
  for (int j = 0; j < 32; ++j) r += __uint_as_float(v[0][j]) + __uint_as_float(v[1][j]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(slot, 512);
}

// mode 6: tcgen05.ld only (two x32 loads in flight, maxima taken so the data is used)
__global__ void kl(float* out, int iters, long long* clk) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(&slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = slot + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 64;
  uint32_t v[2][32];
  float m = 0.f;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; it += 2) {
    tmem_ld32(tb, v[0]);
    tmem_ld32(tb + 32, v[1]);
    tmem_ld_wait_dep(v[0]);
    tmem_ld_wait_dep(v[1]);
    m = fmaxf(m, fmaxf(__uint_as_float(v[0][it & 31]), __uint_as_float(v[1][it & 31])));
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = m;
  if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(slot, 512);
}
void runl(int warps_per_smsp, float* out, long long* clk) {
  const int iters = 2000;
  for (int rep = 0; rep < 2; ++rep) {
    kl<<<148, warps_per_smsp * 128>>>(out, iters, clk);
    cudaDeviceSynchronize();
  }
  long long c;
  cudaMemcpy(&c, clk, 8, cudaMemcpyDeviceToHost);
  printf("mode 6  warps/SMSP %d : %.1f clk per 4 KB chunk per warp -> %.1f B/clk per SMSP\n", warps_per_smsp, (double)c / iters,
         4096.0 * warps_per_smsp * iters / (double)c);
}

template <int MODE>
void runt(int warps_per_smsp, float* out, float* in, long long* clk) {
  const int iters = 2000;
  for (int rep = 0; rep < 2; ++rep) {
    kt<MODE><<<148, warps_per_smsp * 128>>>(out, in, iters, clk);
    cudaDeviceSynchronize();
  }
  long long c;
  cudaMemcpy(&c, clk, 8, cudaMemcpyDeviceToHost);
  printf("mode %d  warps/SMSP %d : %.1f clk per chunk per warp, %.1f clk per chunk per SMSP\n", MODE, warps_per_smsp,
         (double)c / iters, (double)c / iters / warps_per_smsp);
}

template <int MODE>
void run(int warps_per_smsp, float* out, float* in, long long* clk) {
  const int iters = 2000;
  k<MODE><<<148, warps_per_smsp * 128>>>(out, in, iters, clk);
  cudaDeviceSynchronize();
  k<MODE><<<148, warps_per_smsp * 128>>>(out, in, iters, clk);
  cudaDeviceSynchronize();
  long long c;
  cudaMemcpy(&c, clk, 8, cudaMemcpyDeviceToHost);
  printf("mode %d  warps/SMSP %d : %.1f clk per chunk per warp, %.1f clk per chunk per SMSP\n", MODE, warps_per_smsp,
         (double)c / iters, (double)c / iters / warps_per_smsp);
}

int main() {
  float *out, *in;
  long long* clk;
  cudaMalloc(&out, 148 * 1024 * 4);
  cudaMalloc(&in, 1026 * 4);
  cudaMalloc(&clk, 8);
  float h[1026];
  for (int i = 0; i < 1024; ++i) h[i] = -0.001f * i;
  h[1024] = 0.25f;
  h[1025] = 0.5f;
  cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
  for (int w = 1; w <= 4; w *= 2) {
    run<0>(w, out, in, clk);
    run<1>(w, out, in, clk);
    run<2>(w, out, in, clk);
    run<3>(w, out, in, clk);
    runt<4>(w, out, in, clk);
    runt<5>(w, out, in, clk);
    runl(w, out, clk);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
