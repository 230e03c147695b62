// Shared helpers for the funasr_b200 kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/funasr_b200.h"

namespace fb200 {

void set_error(const char* fmt, ...);
void count_launch(int n = 1);

// Times one entry point with CUDA events on the caller's stream when fb200_profile_enable(1)
// was called by this thread; `work` is the algorithmic bytes or FLOPs of the call.
enum ProfKind {
  kProfFbank = 0, kProfLfr, kProfPe, kProfLayerNorm, kProfSplit, kProfFsmn, kProfGemm, kProfAttention,
  kProfCifAlpha, kProfCifScan, kProfCifIntegrate, kProfArgmax, kProfCtc, kProfConvSub, kProfGluDw, kProfRelAttention,
  kProfLstm, kProfUsFires, kProfKinds
};
class ProfScope {
 public:
  ProfScope(int kind, void* stream, double work);
  ~ProfScope();
 private:
  bool active_;
  void* stream_;
};

#define FB_CHECK_ARG(cond, ...)          \
  do {                                   \
    if (!(cond)) {                       \
      fb200::set_error(__VA_ARGS__);     \
      return FB200_EINVAL;               \
    }                                    \
  } while (0)

#define FB_CHECK_LAUNCH(name)                                                        \
  do {                                                                               \
    cudaError_t e_ = cudaGetLastError();                                             \
    if (e_ != cudaSuccess) {                                                         \
      fb200::set_error("%s: launch failed: %s", name, cudaGetErrorString(e_));       \
      return FB200_ECUDA;                                                            \
    }                                                                                \
    fb200::count_launch();                                                           \
  } while (0)

// ---- programmatic dependent launch (PDL) ------------------------------------------------------------------------
// The path is ~26 000 short launches per 1024-utterance step: back to back on one stream, each kernel used to start
// only after the previous one had drained and its completion had travelled through the front end.  Kernels launched
// through launch_pdl() carry cudaLaunchAttributeProgrammaticStreamSerialization: the NEXT kernel's CTAs are scheduled
// as soon as every CTA of the current one has executed pdl_trigger() (all of them do so first thing), run their
// prologue (barrier init, tensor-memory allocation, descriptor prefetch, static weight loads) under the current
// kernel's tail, and block in pdl_wait() until the current kernel has completed and its writes are visible.
// RULE: a kernel launched with launch_pdl() executes pdl_wait() in every thread before it reads or writes any memory
// another kernel of the stream may write (only launch parameters and static weights may be touched before it).
// FB200_PDL=0 in the environment launches everything fully serialised (the instructions are then no-ops).
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
bool pdl_enabled();

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, void* stream, int cluster_x,
                              Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[2];
  unsigned n = 0;
  if (cluster_x > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = (unsigned)cluster_x;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  if (pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

constexpr int kWarp = 32;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// x ~= hi + mid + lo with each part a bf16 (round-to-nearest); hi+mid+lo carries 24 mantissa bits.
__device__ __forceinline__ void split3(float x, __nv_bfloat16& hi, __nv_bfloat16& mid, __nv_bfloat16& lo) {
  hi = __float2bfloat16_rn(x);
  float r = x - __bfloat162float(hi);
  mid = __float2bfloat16_rn(r);
  r -= __bfloat162float(mid);
  lo = __float2bfloat16_rn(r);
}

// Store `x` into a split operand row: part p at column p*kp + col (nparts = 1, 2 or 3).
__device__ __forceinline__ void store_split(__nv_bfloat16* row, int kp, int col, float x, int nparts) {
  const __nv_bfloat16 h = __float2bfloat16_rn(x);
  row[col] = h;
  if (nparts > 1) {
    float r = x - __bfloat162float(h);
    const __nv_bfloat16 m = __float2bfloat16_rn(r);
    row[kp + col] = m;
    if (nparts > 2) row[2 * kp + col] = __float2bfloat16_rn(r - __bfloat162float(m));
  }
}

// Four consecutive values -> 8-byte stores into each part (col % 4 == 0); packed conversions.
__device__ __forceinline__ void store_split4(__nv_bfloat16* row, int kp, int col, const float (&y)[4], int nparts) {
  const __nv_bfloat162 h01 = __floats2bfloat162_rn(y[0], y[1]), h23 = __floats2bfloat162_rn(y[2], y[3]);
  uint2 pk;
  pk.x = *reinterpret_cast<const uint32_t*>(&h01);
  pk.y = *reinterpret_cast<const uint32_t*>(&h23);
  *reinterpret_cast<uint2*>(row + col) = pk;
  if (nparts > 1) {
    const float r0 = y[0] - __low2float(h01), r1 = y[1] - __high2float(h01);
    const float r2 = y[2] - __low2float(h23), r3 = y[3] - __high2float(h23);
    const __nv_bfloat162 m01 = __floats2bfloat162_rn(r0, r1), m23 = __floats2bfloat162_rn(r2, r3);
    pk.x = *reinterpret_cast<const uint32_t*>(&m01);
    pk.y = *reinterpret_cast<const uint32_t*>(&m23);
    *reinterpret_cast<uint2*>(row + kp + col) = pk;
    if (nparts > 2) {
      const __nv_bfloat162 l01 = __floats2bfloat162_rn(r0 - __low2float(m01), r1 - __high2float(m01));
      const __nv_bfloat162 l23 = __floats2bfloat162_rn(r2 - __low2float(m23), r3 - __high2float(m23));
      pk.x = *reinterpret_cast<const uint32_t*>(&l01);
      pk.y = *reinterpret_cast<const uint32_t*>(&l23);
      *reinterpret_cast<uint2*>(row + 2 * kp + col) = pk;
    }
  }
}

__host__ __device__ __forceinline__ int round_up(int x, int m) { return (x + m - 1) / m * m; }

// Sequence index of packed row r: largest b with cu[b] <= r (cu has batch+1 entries).
__device__ __forceinline__ int find_seq(const int32_t* __restrict__ cu, int batch, int r) {
  int lo = 0, hi = batch;  // invariant: cu[lo] <= r < cu[hi]
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (__ldg(cu + mid) <= r) lo = mid; else hi = mid;
  }
  return lo;
}

}  // namespace fb200
