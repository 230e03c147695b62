// Rate of tcgen05.mma cta_group::2 (M = 256 over a CTA pair, N = 256): kind::f16 (K = 16 per instruction) against
// kind::f8f6f4 with e4m3 operands (K = 32 per instruction), and a functional check of the scale-input-d immediate
// (D = A*B + D * 2^-s) that lets fp8 correction terms be accumulated at a different scale than the fp16 main term.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I funasr_b200/csrc tools/micro/mma_fp8_rate.cu -o tools/micro/mma_fp8_rate
#include <cstdio>
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include "common.cuh"
#include "sm100_ptx.cuh"
namespace fb200 { void set_error(const char*, ...) {} }
using namespace fb200;

__host__ __device__ constexpr uint32_t idesc_f16(int m, int n) {  // fp16 x fp16 -> fp32
  return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
__host__ __device__ constexpr uint32_t idesc_e4m3(int m, int n) {  // e4m3 x e4m3 -> fp32 (kind::f8f6f4)
  return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
template <int KIND>
__device__ __forceinline__ void mma_pair(uint32_t d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  if (KIND == 0)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                 "l"(da), "l"(db), "r"(idesc), "r"(acc)
                 : "memory");
  else
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                 "l"(da), "l"(db), "r"(idesc), "r"(acc)
                 : "memory");
}

template <int KIND>
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
    constexpr uint32_t idesc = KIND == 0 ? idesc_f16(256, 256) : idesc_e4m3(256, 256);
    for (int r = 0; r < reps; r += 16) {
#pragma unroll
      for (int s = 0; s < 16; ++s) {
        const uint64_t da = umma_desc_sw128(base + (s & 3) * 32 + (s >> 2) * 16384);
        const uint64_t db = umma_desc_sw128(base + 65536 + (s & 3) * 32 + (s >> 2) * 16384);
        mma_pair<KIND>(tmem, da, db, idesc, 1u);
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

// One CTA, M = 128, N = 64: D = (e4m3 A8 * B8) then D = fp16 A16 * B16 + D * 2^-4; A, B rows constant -> D known.
__global__ void __launch_bounds__(128, 1) scale_kernel(float* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t base = smem_u32(smem);
  const uint32_t bar = base + 65536, slot = bar + 8;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    if (elect_one()) { mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "r"(64u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  // A8 [128 x 128 B] at 0: e4m3 1.0 = 0x38 everywhere; B8 [64 x 128 B] at 16384: e4m3 2.0 = 0x40
  // A16 [128 x 64 fp16] at 32768: 0.5; B16 [64 x 64 fp16] at 49152: 0.25      (constant tiles: swizzle-invariant)
  for (int i = threadIdx.x; i < 16384; i += blockDim.x) smem[i] = 0x38;
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) smem[16384 + i] = 0x40;
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) reinterpret_cast<__half*>(smem + 32768)[i] = __float2half(0.5f);
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) reinterpret_cast<__half*>(smem + 49152)[i] = __float2half(0.25f);
  fence_proxy_async_smem();
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem) : "r"(slot));
  if (warp == 0 && elect_one()) {
    const uint32_t i8 = idesc_e4m3(128, 64), i16 = idesc_f16(128, 64);
    for (int s = 0; s < 4; ++s) {  // K = 128 of fp8: 4 x 32 -> D = 128 * 1 * 2 = 256
      const uint64_t da = umma_desc_sw128(base + s * 32), db = umma_desc_sw128(base + 16384 + s * 32);
      asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem),
                   "l"(da), "l"(db), "r"(i8), "r"(s > 0 ? 1u : 0u)
                   : "memory");
    }
    for (int s = 0; s < 4; ++s) {  // K = 64 of fp16: 4 x 16 -> + 64 * 0.5 * 0.25 = 8; first one scales D by 2^-4 -> 16 + 8 = 24
      const uint64_t da = umma_desc_sw128(base + 32768 + s * 32), db = umma_desc_sw128(base + 49152 + s * 32);
      if (s == 0)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p, 4;\n\t}" ::"r"(tmem),
                     "l"(da), "l"(db), "r"(i16)
                     : "memory");
      else
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem),
                     "l"(da), "l"(db), "r"(i16)
                     : "memory");
    }
    umma_commit(bar);
  }
  mbar_wait(bar, 0);
  tcgen05_fence_after();
  uint32_t v[32];
  tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16), v);
  if (threadIdx.x == 0 || threadIdx.x == 127) { out[threadIdx.x ? 2 : 0] = __uint_as_float(v[0]); out[threadIdx.x ? 3 : 1] = __uint_as_float(v[31]); }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64u));
  (void)lane;
}

template <int KIND>
void run(int grid, long long* d) {
  const int reps = 512;
  cudaFuncSetAttribute(rate_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200000);
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
    cudaLaunchKernelEx(&cfg, rate_kernel<KIND>, reps, d);
    cudaDeviceSynchronize();
  }
  cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  printf("grid %3d pair SS M=256 N=256 %s: issue %.1f cyc/mma, complete %.1f cyc/mma  err=%s\n", grid,
         KIND == 0 ? "kind::f16 K=16   " : "kind::f8f6f4 K=32", (double)h[0] / reps, (double)h[1] / reps,
         cudaGetErrorString(cudaGetLastError()));
}

int main() {
  long long* d;
  cudaMalloc(&d, 16);
  for (int grid : {2, 148}) {
    run<0>(grid, d);
    run<1>(grid, d);
  }
  float* f;
  cudaMalloc(&f, 16);
  cudaMemset(f, 0, 16);
  cudaFuncSetAttribute(scale_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 70000);
  scale_kernel<<<1, 128, 70000>>>(f);
  cudaDeviceSynchronize();
  float hf[4];
  cudaMemcpy(hf, f, 16, cudaMemcpyDeviceToHost);
  printf("scale-input-d: D = %g %g %g %g (expect 24: (128*1*2) * 2^-4 + 64*0.5*0.25)  err=%s\n", hf[0], hf[1], hf[2], hf[3],
         cudaGetErrorString(cudaGetLastError()));
  return 0;
}
