// Issue-to-completion rate of tcgen05.mma kind::f16 (M = 128) for several N, with A from shared memory (SS) or
// tensor memory (TS).  Standalone: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I funasr_b200/csrc
// tools/micro/mma_rate.cu -o mma_rate
#include <cstdio>
#include <cuda.h>
#include <cuda_bf16.h>
#include "common.cuh"
#include "sm100_ptx.cuh"
namespace fb200 { void set_error(const char*, ...) {} }
using namespace fb200;

// SPIN > 0: that many other warps poll an mbarrier that never completes while warp 0 issues (what the waiting roles of a
// warp-specialised kernel do)
template <int N, int TS, int SPIN = 0>
__global__ void __launch_bounds__(256, 1) rate_kernel(int reps, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t base = smem_u32(smem);
  const uint32_t bar = base + 196608, slot = bar + 8;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    if (elect_one()) { mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "r"(512u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  for (int i = threadIdx.x; i < 196608 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  fence_proxy_async_smem();
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem) : "r"(slot));
  __shared__ volatile int done;
  if (threadIdx.x == 0) { done = 0; mbar_init(bar + 16, 1); }
  __syncthreads();
  if (warp >= 1 && warp <= SPIN) {
    while (!done) mbar_try_wait(bar + 16, 0);
  }
  if (warp == 0 && elect_one()) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, N);
    long long t0 = clock64();
    for (int r = 0; r < reps; r += 16) {
#pragma unroll
      for (int s = 0; s < 16; ++s) {  // operands rotate over 16 K-slices so that shared-memory reads are real
        const uint64_t da = umma_desc_sw128(base + (s & 3) * 32 + (s >> 2) * 16384);
        const uint64_t db = umma_desc_sw128(base + 65536 + (s & 3) * 32 + (s >> 2) * 32768);
        if (TS)
          umma_bf16_ts(tmem, tmem + 256 + (s & 7) * 8, db, idesc, 1u);
        else
          umma_bf16(tmem, da, db, idesc, 1u);
      }
    }
    long long t1 = clock64();
    umma_commit(bar);
    mbar_wait(bar, 0);
    long long t2 = clock64();
    if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    done = 1;
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u));
}

template <int N, int TS, int SPIN = 0>
void run(int grid, long long* d) {
  const int reps = 512;
  cudaFuncSetAttribute(rate_kernel<N, TS, SPIN>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200000);
  long long h[2];
  for (int it = 0; it < 2; ++it) {
    rate_kernel<N, TS, SPIN><<<grid, 256, 200000>>>(reps, d);
    cudaDeviceSynchronize();
  }
  cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  printf("spin %d grid %3d %s N=%3d: issue %.1f cyc/mma, complete %.1f cyc/mma (nominal %d)  err=%s\n", SPIN, grid, TS ? "TS" : "SS", N,
         (double)h[0] / reps, (double)h[1] / reps, N / 2, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  long long* d;
  cudaMalloc(&d, 16);
  run<32, 0, 1>(148, d);
  run<32, 0, 3>(148, d);
  run<32, 0, 6>(148, d);
  run<64, 0, 6>(148, d);
  for (int grid : {148}) {
    run<16, 0>(grid, d);
    run<32, 0>(grid, d);
    run<32, 1>(grid, d);
    run<64, 0>(grid, d);
    run<128, 0>(grid, d);
    run<256, 0>(grid, d);
    run<64, 1>(grid, d);
    run<128, 1>(grid, d);
    run<256, 1>(grid, d);
  }
  return 0;
}
