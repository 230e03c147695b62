// Issue-to-completion rate of tcgen05.mma cta_group::2 kind::f16 (M = 256 over a CTA pair), A and B from shared memory.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I funasr_b200/csrc tools/micro/mma_pair_rate.cu -o mma_pair_rate
#include <cstdio>
#include <cuda.h>
#include <cuda_bf16.h>
#include "common.cuh"
#include "sm100_ptx.cuh"
namespace fb200 { void set_error(const char*, ...) {} }
using namespace fb200;

template <int N>
__global__ void __launch_bounds__(128, 1) rate_kernel(int reps, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t base = smem_u32(smem);
  const uint32_t bar = base + 196608, slot = bar + 8;
  const int warp = threadIdx.x >> 5;
  const uint32_t rank = cluster_ctarank();
  if (warp == 0) {
    if (elect_one()) { mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "r"(512u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::);
  }
  for (int i = threadIdx.x; i < 196608 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  fence_proxy_async_smem();
  tcgen05_fence_before();
  cluster_sync_all();
  tcgen05_fence_after();
  uint32_t tmem;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem) : "r"(slot));
  long long t0 = clock64();
  if (warp == 0 && rank == 0 && elect_one()) {
    constexpr uint32_t idesc = umma_idesc_bf16(256, N);
    for (int r = 0; r < reps; r += 16) {
#pragma unroll
      for (int s = 0; s < 16; ++s) {
        const uint64_t da = umma_desc_sw128(base + (s & 3) * 32 + (s >> 2) * 16384);
        const uint64_t db = umma_desc_sw128(base + 65536 + (s & 3) * 32 + (s >> 2) * 16384);
        umma_bf16_pair(tmem, da, db, idesc, 1u);
      }
    }
    long long t1 = clock64();
    umma_commit_pair(bar, 3);
    if (blockIdx.x == 0) out[0] = t1 - t0;
  }
  if (warp == 0) {
    mbar_wait(bar, 0);
    long long t2 = clock64();
    if (blockIdx.x == 0 && (threadIdx.x & 31) == 0) out[1] = t2 - t0;
  }
  tcgen05_fence_before();
  cluster_sync_all();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u));
}

template <int N>
void run(int grid, long long* d) {
  const int reps = 512;
  cudaFuncSetAttribute(rate_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200000);
  long long h[2];
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = 200000;
  cudaLaunchAttribute at;
  at.id = cudaLaunchAttributeClusterDimension;
  at.val.clusterDim.x = 2;
  at.val.clusterDim.y = at.val.clusterDim.z = 1;
  cfg.attrs = &at;
  cfg.numAttrs = 1;
  for (int it = 0; it < 2; ++it) {
    cudaLaunchKernelEx(&cfg, rate_kernel<N>, reps, d);
    cudaDeviceSynchronize();
  }
  cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  printf("grid %3d pair SS M=256 N=%3d: issue %.1f cyc/mma, complete %.1f cyc/mma (nominal %d)  err=%s\n", grid, N,
         (double)h[0] / reps, (double)h[1] / reps, N / 2, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  long long* d;
  cudaMalloc(&d, 16);
  for (int grid : {2, 148}) {
    run<64>(grid, d);
    run<128>(grid, d);
    run<256>(grid, d);
  }
  return 0;
}
