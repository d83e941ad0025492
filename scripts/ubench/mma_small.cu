// Micro-benchmark: issue-to-completion time of a run of small tcgen05 MMAs (M128 x N x K16, bf16, A from
// TMEM, B from shared memory) accumulating into ONE or into several round-robin accumulators.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -o mma_small mma_small.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../copilot-for-consensus_b200/csrc/common.cuh"
using namespace b200;

template <int N, int NACC, bool TS>
__global__ void k(long long* out, int nmma) {
  __shared__ __align__(1024) uint8_t smem[32768];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 32768 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&slot, 512);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = slot;
  if (warp == 1 && lane == 0) {
    const uint32_t idesc = TS ? umma_idesc_bf16_f32_bmn(128, N) : umma_idesc_bf16_f32(128, N);
    const uint32_t sb = smem_u32(smem);
    for (int rep = 0; rep < 3; ++rep) {
      const long long t0 = clock64();
      for (int i = 0; i < nmma; ++i) {
        const uint32_t d = tb + 256 + (i % NACC) * N;
        if (TS) {
          const uint64_t db = umma_smem_desc(sb + (i & 15) * 1024, 4, 512, 16384);
          umma_bf16_ts(d, tb + (i & 15) * 8, db, idesc, (uint32_t)(i >= NACC));
        } else {
          const uint64_t da = umma_smem_desc(sb + 16384 + (i & 1) * 32, 4, 512, 0);
          const uint64_t db = umma_smem_desc(sb + (i & 1) * 32, 4, 512, 0);
          umma_bf16(d, da, db, idesc, (uint32_t)(i >= NACC));
        }
      }
      const long long t1 = clock64();
      umma_commit(&bar);
      mbar_wait(&bar, rep & 1);
      const long long t2 = clock64();
      if (blockIdx.x == 0) {
        out[rep * 2] = t1 - t0;
        out[rep * 2 + 1] = t2 - t0;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tb, 512);
}

template <int N, int NACC, bool TS>
void run(long long* d) {
  const int nmma = 64;
  k<N, NACC, TS><<<148, 128>>>(d, nmma);
  cudaDeviceSynchronize();
  long long h[6];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  printf("%s N=%3d acc=%d : issue %6.1f clk/MMA, complete %6.1f clk/MMA   (%s)\n", TS ? "TS" : "SS", N, NACC,
         (double)h[4] / nmma, (double)h[5] / nmma, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  long long* d;
  cudaMalloc(&d, 64);
  run<32, 1, true>(d);
  run<32, 2, true>(d);
  run<32, 4, true>(d);
  run<64, 1, true>(d);
  run<64, 2, true>(d);
  run<32, 1, false>(d);
  run<32, 4, false>(d);
  run<64, 1, false>(d);
  run<128, 1, false>(d);
  run<256, 1, false>(d);
  run<192, 1, false>(d);
  run<160, 1, false>(d);
  run<96, 1, false>(d);
  run<224, 1, false>(d);
  return 0;
}
