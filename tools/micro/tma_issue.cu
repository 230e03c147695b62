// TMA load issue rate / latency from ONE thread: N loads of a [rows x 64] bf16 box (128-byte swizzle) into a shared-memory
// ring, descriptor passed as a __grid_constant__ kernel parameter or read from global memory, one or `grid` CTAs, every CTA
// reading the SAME tile or its own.  Prints cycles to issue all loads and until the last byte landed.
// Standalone: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I funasr_b200/csrc tools/micro/tma_issue.cu -o tma_issue -lcuda
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <cuda_bf16.h>
#include "common.cuh"
#include "sm100_ptx.cuh"
namespace fb200 { void set_error(const char*, ...) {} }
using namespace fb200;

__global__ void __launch_bounds__(128, 1)
issue_kernel(const __grid_constant__ CUtensorMap pmap, const CUtensorMap* gmap, int use_global, int loads, int box_bytes,
             int same_tile, int rows_total, int box_rows, int nwarps, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t base = (smem_u32(smem) + 1023u) & ~1023u;
  const uint32_t bar = base + 196608;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) mbar_expect_tx(bar, (uint32_t)loads * (uint32_t)box_bytes);
  __syncthreads();
  if ((threadIdx.x & 31) == 0 && (threadIdx.x >> 5) < nwarps) {  // lane 0 of the first `nwarps` warps share the loads
    const CUtensorMap* map = use_global ? gmap : &pmap;
    const int slots = 196608 / box_bytes;
    const int row_tiles = rows_total / box_rows;
    long long t0 = clock64();
    for (int i = threadIdx.x >> 5; i < loads; i += nwarps) {
      const int tile = same_tile ? i : (i + blockIdx.x * 7);
      tma_load_2d(base + (uint32_t)(i % slots) * box_bytes, map, (tile & 7) * 64, ((tile >> 3) % row_tiles) * box_rows, bar);
    }
    long long t1 = clock64();
    mbar_wait(bar, 0);
    long long t2 = clock64();
    if (blockIdx.x == 0 && threadIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
}

static CUtensorMap make(void* base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
  typedef CUresult (*Fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                         const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* ptr = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q);
  CUtensorMap m;
  cuuint64_t dims[2] = {cols, rows}, strides[1] = {cols * 2};
  cuuint32_t box[2] = {64, box_rows}, es[2] = {1, 1};
  CUresult r = ((Fn)ptr)(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(1); }
  return m;
}

int main() {
  const uint64_t rows = 4096, cols = 512;
  void* buf;
  cudaMalloc(&buf, rows * cols * 2);
  cudaMemset(buf, 0, rows * cols * 2);
  long long* d;
  cudaMalloc(&d, 16);
  cudaFuncSetAttribute(issue_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200000);
  for (int box_rows : {32, 128}) {
    CUtensorMap m = make(buf, rows, cols, box_rows);
    CUtensorMap* gm;
    cudaMalloc(&gm, sizeof(m));
    cudaMemcpy(gm, &m, sizeof(m), cudaMemcpyHostToDevice);
    for (int use_global : {1})
      for (int grid : {148})
        for (int nwarps : {1, 2, 4})
        for (int same : {0}) {
          const int loads = 48, box_bytes = box_rows * 128;
          long long h[2] = {0, 0};
          for (int rep = 0; rep < 3; ++rep) {  // third run: warm L2 / descriptor
            issue_kernel<<<grid, 128, 200000>>>(m, gm, use_global, loads, box_bytes, same, (int)rows, box_rows, nwarps, d);
            cudaDeviceSynchronize();
          }
          cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
          printf("warps %d  box %3d rows (%5d B)  map %-6s grid %3d  %s tile: issue %6lld cyc (%4lld / load), landed %6lld cyc -> %5.1f B/clk\n", nwarps, box_rows,
                 box_bytes, use_global ? "global" : "param", grid, same ? "same" : "own ", h[0], h[0] / loads, h[1],
                 (double)loads * box_bytes / (double)h[1]);
        }
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
