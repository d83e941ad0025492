// Micro-benchmark: do warps saturating the SFU (MUFU.EX2) slow down OTHER warps on the same scheduler
// that only issue FMA-pipe work?  Warps 0-7 (two per scheduler) run the softmax inner loop, warps 8-15
// run independent FFMA chains; each kind reports its own clocks with and without the other present.
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

__global__ void k(float* out, const float* in, int iters, long long* clk, int run_sfu, int run_fma) {
  const int warp = threadIdx.x >> 5;
  const bool sfu = warp < 8;
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = in[(threadIdx.x * 32 + j) & 1023];
  const float scale = in[1024];
  float ms = in[1025], s0 = 0.f, s1 = 0.f;
  __syncthreads();
  const long long t0 = clock64();
  if (sfu && run_sfu) {
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
      ms += 1e-4f;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        s0 += ex2(fmaf(v[2 * j], scale, -ms));
        s1 += ex2(fmaf(v[2 * j + 1], scale, -ms));
      }
    }
  } else if (!sfu && run_fma) {
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = fmaf(v[j], scale, ms);   // 128 independent-ish FFMA per iteration
    }
  }
  const long long t1 = clock64();
  float r = s0 + s1;
#pragma unroll
  for (int j = 0; j < 32; ++j) r += v[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (blockIdx.x == 0 && (threadIdx.x == 0 || threadIdx.x == 256)) clk[threadIdx.x ? 1 : 0] = t1 - t0;
}

int main() {
  float *out, *in; long long* clk;
  cudaMalloc(&out, 148 * 512 * 4); cudaMalloc(&in, 1026 * 4); cudaMalloc(&clk, 16);
  float h[1026]; for (int i = 0; i < 1024; ++i) h[i] = -0.001f * i; h[1024] = 0.25f; h[1025] = 0.5f;
  cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
  const int iters = 2000;
  for (int mode = 0; mode < 3; ++mode) {
    const int rs = mode != 1, rf = mode != 0;
    for (int rep = 0; rep < 2; ++rep) { k<<<148, 512>>>(out, in, iters, clk, rs, rf); cudaDeviceSynchronize(); }
    long long c[2]; cudaMemcpy(c, clk, 16, cudaMemcpyDeviceToHost);
    printf("sfu warps %s, fma warps %s : sfu %.1f clk/iter (32 MUFU), fma %.1f clk/iter (128 FFMA)\n", rs ? "on " : "off", rf ? "on " : "off",
           (double)c[0] / iters, (double)c[1] / iters);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
