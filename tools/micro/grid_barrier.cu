// Device-wide barrier variants on 148 resident CTAs x 256 threads: ns per barrier, back to back, no work in between.
//   0  __syncthreads; thread 0: red.release.gpu + poll ld.acquire.gpu; __syncthreads            (stack kernel, first version)
//   1  every warp's lane 0 fences (membar.gl) BEFORE __syncthreads; thread 0: relaxed red + poll ld.relaxed + fence after
//   2  like 1, polled by all 32 lanes of warp 0 (volatile loads in flight back to back)
//   3  like 1 with atom.add (the arrival returns the count: the last CTA skips the poll)
// Standalone: nvcc -gencode arch=compute_100a,code=sm_100a -O3 tools/micro/grid_barrier.cu -o grid_barrier
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned ld_acquire(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned ld_relaxed(const unsigned* p) {
  unsigned v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

template <int V>
__global__ void __launch_bounds__(256, 1) bar_kernel(unsigned* counter, int reps, float* sink, long long* out) {
  unsigned gen = 0;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  long long t0 = 0;
  for (int r = 0; r < reps; ++r) {
    if (r == 8 && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    sink[blockIdx.x * 256 + threadIdx.x] = (float)r;  // a global write the barrier has to publish
    ++gen;
    const unsigned target = gen * gridDim.x;
    if (V == 0) {
      __syncthreads();
      if (threadIdx.x == 0) {
        asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(counter), "r"(1u) : "memory");
        while (ld_acquire(counter) < target) {}
      }
      __syncthreads();
    } else if (V == 1 || V == 3) {
      if (lane == 0) __threadfence();
      __syncthreads();
      if (threadIdx.x == 0) {
        bool last = false;
        if (V == 1) asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(counter), "r"(1u) : "memory");
        else last = atomicAdd(counter, 1u) + 1 == target;
        if (!last) while (ld_relaxed(counter) < target) {}
        __threadfence();
      }
      __syncthreads();
    } else {
      if (lane == 0) __threadfence();
      __syncthreads();
      if (warp == 0) {
        if (lane == 0) asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(counter), "r"(1u) : "memory");
        bool seen = false;
        while (!seen) {
          seen = __any_sync(0xffffffffu, ld_relaxed(counter) >= target);
        }
        if (lane == 0) __threadfence();
      }
      __syncthreads();
    }
  }
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    out[0] = t1 - t0;
  }
}

template <int V>
void run(unsigned* c, float* sink, long long* d) {
  const int reps = 1008;
  cudaMemset(c, 0, 4);
  void* args[] = {&c, (void*)&reps, &sink, &d};
  cudaLaunchCooperativeKernel((const void*)bar_kernel<V>, dim3(148), dim3(256), args, 0, 0);
  cudaDeviceSynchronize();
  long long h = 0;
  cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
  printf("variant %d: %.0f ns per barrier  (%s)\n", V, (double)h / (reps - 8), cudaGetErrorString(cudaGetLastError()));
}

int main() {
  unsigned* c;
  float* sink;
  long long* d;
  cudaMalloc(&c, 256);
  cudaMalloc(&sink, 148 * 256 * 4);
  cudaMalloc(&d, 16);
  for (int k = 0; k < 2; ++k) {
    run<0>(c, sink, d);
    run<1>(c, sink, d);
    run<2>(c, sink, d);
    run<3>(c, sink, d);
  }
  return 0;
}
