// Does sm_100a have a packed-half exponential on the SFU, and at what rate?  Per warp: cycles per 32-score chunk.
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t ex2_h2(uint32_t x) { uint32_t y; asm volatile("ex2.approx.f16x2 %0, %1;" : "=r"(y) : "r"(x)); return y; }
__device__ __forceinline__ uint32_t ex2_bf2(uint32_t x) { uint32_t y; asm volatile("ex2.approx.ftz.bf16x2 %0, %1;" : "=r"(y) : "r"(x)); return y; }
template <int MODE>
__global__ void k(float* out, const float* in, int iters, long long* clk) {
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = in[(threadIdx.x * 32 + j) & 1023];
  uint32_t acc = 0;
  float ms = in[1025];
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; ++it) {
    ms += 1e-4f;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const float a0 = v[2 * j] - ms, a1 = v[2 * j + 1] - ms;
      if (MODE == 0) {                       // fp32 MUFU x2 + pack to bf16x2
        __nv_bfloat162 t = __floats2bfloat162_rn(ex2(a0), ex2(a1));
        acc ^= *reinterpret_cast<uint32_t*>(&t);
      } else if (MODE == 1) {                // pack to half2, one packed MUFU
        __half2 h = __floats2half2_rn(a0, a1);
        acc ^= ex2_h2(*reinterpret_cast<uint32_t*>(&h));
      } else {                               // pack to bf16x2, one packed MUFU
        __nv_bfloat162 h = __floats2bfloat162_rn(a0, a1);
        acc ^= ex2_bf2(*reinterpret_cast<uint32_t*>(&h));
      }
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = __uint_as_float(acc) + ms;
  if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}
int main() {
  float *in, *out; long long* clk;
  cudaMalloc(&in, 4096 * 4); cudaMalloc(&out, 1 << 20); cudaMalloc(&clk, 8);
  cudaMemset(in, 0, 4096 * 4);
  const int iters = 2000;
  for (int warps = 1; warps <= 4; warps *= 2) {
    long long c[3];
    k<0><<<1, 128 * warps, 0>>>(out, in, iters, clk); cudaMemcpy(&c[0], clk, 8, cudaMemcpyDeviceToHost);
    k<1><<<1, 128 * warps, 0>>>(out, in, iters, clk); cudaMemcpy(&c[1], clk, 8, cudaMemcpyDeviceToHost);
    k<2><<<1, 128 * warps, 0>>>(out, in, iters, clk); cudaMemcpy(&c[2], clk, 8, cudaMemcpyDeviceToHost);
    printf("warps/scheduler %d: clocks per 32-score chunk per warp: f32 %.1f  f16x2 %.1f  bf16x2 %.1f  (err %s)\n", warps,
           (double)c[0] / iters / warps, (double)c[1] / iters / warps, (double)c[2] / iters / warps, cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
