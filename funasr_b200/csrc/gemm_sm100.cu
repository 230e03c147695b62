// Dense contraction C = A . W^T over bf16 split operands on the 5th-gen tensor cores:
//   TMA (cp.async.bulk.tensor, 128B swizzle) -> shared memory ring -> tcgen05.mma (one elected
//   thread, fp32 accumulators in TMEM) -> tcgen05.ld -> fused epilogue (bias / ReLU / two
//   residual addends / fp32 and bf16-split outputs).
// fp32-parity mode runs the 6 cross terms of the 3-way bf16 split as extra K iterations of the
// same accumulator (small terms first), so no separate accumulation pass exists.
// A SIMT fp32 kernel over the SAME operands and epilogue is kept for bring-up tests (impl=1).
#include <cuda.h>

#include "common.cuh"

namespace fb200 {

constexpr int BM = 128;
constexpr int BK = 64;  // bf16 elements = 128 bytes = one swizzle row
constexpr int kStages = 4;
constexpr int kGemmThreads = 192;  // warp0 TMA, warp1 MMA + TMEM alloc, warps 2..5 epilogue

struct GemmParams {
  int m, n, k_blocks, kp_a, kp_w;
  int taps, nterms;
  int term_a[6], term_b[6];
  const float* bias;
  int relu;
  const float* res1;
  int ld_res1;
  const float* res2;
  int ld_res2;
  float* out_f32;
  int ld_out;
  __nv_bfloat16* out_split;
  int out_parts, out_kp;
};

// ---------------------------------------------------------------- PTX wrappers ----------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok;
}
// Bounded wait: a pipeline bug must trap (reported as a CUDA error), never hang the GPU box.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000ll) {
      printf("fb200 gemm: mbarrier wait timed out (block %d,%d thread %d bar %u parity %u)\n", blockIdx.x, blockIdx.y,
             threadIdx.x, bar, parity);
      __trap();
    }
  }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// K-major operand tile, 128-byte swizzle: rows are 128 B, 8-row atoms are 1024 B apart.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);  // start address
  d |= (uint64_t)1 << 16;                        // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;              // stride byte offset
  d |= (uint64_t)1 << 46;                        // descriptor version (sm_100)
  d |= (uint64_t)2 << 61;                        // SWIZZLE_128B
  return d;
}
// kind::f16 instruction descriptor: bf16 x bf16 -> fp32, both operands K-major.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int m, int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// ---------------------------------------------------------------- shared epilogue -------
// 32 consecutive output columns [col0, col0+32) of row `row`.
__device__ __forceinline__ void epilogue_store32(const GemmParams& p, int row, int col0, float* acc) {
  if (row >= p.m) return;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    const int col = col0 + j;
    if (col < p.n) {
      float y = acc[j];
      if (p.bias) y += __ldg(p.bias + col);
      if (p.relu) y = fmaxf(y, 0.f);
      if (p.res1) y += p.res1[(size_t)row * p.ld_res1 + col];
      if (p.res2) y += p.res2[(size_t)row * p.ld_res2 + col];
      acc[j] = y;
    } else {
      acc[j] = 0.f;
    }
  }
  if (p.out_f32) {
    float* o = p.out_f32 + (size_t)row * p.ld_out + col0;
    if (col0 + 32 <= p.n && (p.ld_out & 3) == 0) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(acc[j], acc[j + 1], acc[j + 2], acc[j + 3]);
    } else {
      for (int j = 0; j < 32; ++j)
        if (col0 + j < p.n) o[j] = acc[j];
    }
  }
  if (p.out_split) {
    // padding columns n..out_kp are written as zero (acc was zeroed above)
    __nv_bfloat16* o = p.out_split + (size_t)row * (p.out_parts * p.out_kp);
    if (col0 + 32 <= p.out_kp) {
      __nv_bfloat16 h[32], mi[32], lo[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        if (p.out_parts == 1) h[j] = __float2bfloat16_rn(acc[j]);
        else split3(acc[j], h[j], mi[j], lo[j]);
      }
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        *reinterpret_cast<uint4*>(o + col0 + j) = *reinterpret_cast<uint4*>(h + j);
        if (p.out_parts == 3) {
          *reinterpret_cast<uint4*>(o + p.out_kp + col0 + j) = *reinterpret_cast<uint4*>(mi + j);
          *reinterpret_cast<uint4*>(o + 2 * p.out_kp + col0 + j) = *reinterpret_cast<uint4*>(lo + j);
        }
      }
    } else {
      for (int j = 0; j < 32; ++j)
        if (col0 + j < p.out_kp) store_split(o, p.out_kp, col0 + j, acc[j], p.out_parts);
    }
  }
}

// ---------------------------------------------------------------- tcgen05 kernel --------
template <int BN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                    const GemmParams p) {
  constexpr uint32_t kABytes = BM * BK * 2;
  constexpr uint32_t kBBytes = BN * BK * 2;
  constexpr uint32_t kStageBytes = kABytes + kBBytes;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B needs 1024 B alignment
  const uint32_t bar_base = smem_base + kStages * kStageBytes;
  // barriers: full[kStages], empty[kStages], tmem_full; then the TMEM base address slot
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (2 * kStages);
  const uint32_t tmem_slot = bar_base + 8u * (2 * kStages + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * BN, m0 = blockIdx.y * BM;
  const int iters = p.taps * p.nterms * p.k_blocks;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < kStages; ++s) {
        mbar_init(full_bar(s), 1);
        mbar_init(empty_bar(s), 1);
      }
      mbar_init(tmem_full_bar, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"((uint32_t)BN));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp == 0) {
    if (lane == 0) {  // ===== TMA producer =====
      int it = 0;
      for (int tap = 0; tap < p.taps; ++tap)
        for (int t = 0; t < p.nterms; ++t)
          for (int kb = 0; kb < p.k_blocks; ++kb, ++it) {
            const int s = it % kStages;
            const uint32_t ph = (uint32_t)(it / kStages) & 1u;
            mbar_wait(empty_bar(s), ph ^ 1u);
            mbar_expect_tx(full_bar(s), kStageBytes);
            const uint32_t sa = smem_base + s * kStageBytes;
            tma_load_2d(sa, &map_a, p.term_a[t] * p.kp_a + kb * BK, m0 + tap, full_bar(s));
            tma_load_2d(sa + kABytes, &map_w, p.term_b[t] * p.kp_w + kb * BK, tap * p.n + n0, full_bar(s));
          }
    }
  } else if (warp == 1) {
    if (lane == 0) {  // ===== MMA issuer =====
      constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
      for (int it = 0; it < iters; ++it) {
        const int s = it % kStages;
        const uint32_t ph = (uint32_t)(it / kStages) & 1u;
        mbar_wait(full_bar(s), ph);
        tcgen05_fence_after();
        const uint32_t sa = smem_base + s * kStageBytes;
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          const uint64_t da = umma_desc_sw128(sa + k * 32);
          const uint64_t db = umma_desc_sw128(sa + kABytes + k * 32);
          umma_bf16(tmem_base, da, db, idesc, (it > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(empty_bar(s));  // frees the smem slot once these MMAs retire
      }
      umma_commit(tmem_full_bar);
    }
  } else {  // ===== epilogue warps 2..5: TMEM lane group = warp % 4 =====
    const int lg = warp & 3;
    mbar_wait(tmem_full_bar, 0);
    tcgen05_fence_after();
    const int row = m0 + lg * 32 + lane;
#pragma unroll 1
    for (int c = 0; c < BN; c += 32) {
      uint32_t r[32];
      tmem_ld32(tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)c, r);
      if (n0 + c < p.n || (p.out_split && n0 + c < p.out_kp))
        epilogue_store32(p, row, n0 + c, reinterpret_cast<float*>(r));
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)BN));
  }
}

// ---------------------------------------------------------------- SIMT check kernel -----
// Same operands, same term list, same epilogue; fp32 FFMA.  One thread = one row x 32 columns.
__global__ void gemm_simt_kernel(const __nv_bfloat16* __restrict__ a, int a_rows, int lda,
                                 const __nv_bfloat16* __restrict__ w, int ldw, const GemmParams p) {
  const int row = blockIdx.y * blockDim.x + threadIdx.x;
  const int col0 = blockIdx.x * 32;
  float acc[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) acc[j] = 0.f;
  if (row < p.m) {
    for (int tap = 0; tap < p.taps; ++tap) {
      const int ar = row + tap;
      if (ar >= a_rows) continue;
      for (int t = 0; t < p.nterms; ++t) {
        const __nv_bfloat16* ap = a + (size_t)ar * lda + p.term_a[t] * p.kp_a;
        for (int k = 0; k < p.k_blocks * BK; ++k) {
          const float av = __bfloat162float(ap[k]);
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int col = col0 + j;
            if (col < p.n)
              acc[j] = fmaf(av, __bfloat162float(w[((size_t)tap * p.n + col) * ldw + p.term_b[t] * p.kp_w + k]), acc[j]);
          }
        }
      }
    }
  }
  epilogue_store32(p, row, col0, acc);
}

// ---------------------------------------------------------------- host side -------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)ptr;
  }
  return fn;
}

// 2-D bf16 row-major [rows, cols] tensor, box = [box_rows, 64 cols], 128-byte swizzle, zero OOB fill.
static int make_map(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (fn == nullptr) {
    set_error("gemm: cuTensorMapEncodeTiled entry point not available");
    return FB200_ECUDA;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("gemm: cuTensorMapEncodeTiled failed (%d) rows=%llu cols=%llu", (int)r, (unsigned long long)rows,
              (unsigned long long)cols);
    return FB200_ECUDA;
  }
  return FB200_OK;
}

template <int BN>
static int launch_tcgen05(const CUtensorMap& ma, const CUtensorMap& mw, const GemmParams& p, cudaStream_t st) {
  constexpr size_t smem = kStages * (BM * BK * 2 + BN * BK * 2) + 1024 + 256;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tcgen05_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
      set_error("gemm: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return FB200_ECUDA;
    }
    attr_set = true;
  }
  const int ncols = p.out_split ? p.out_kp : p.n;  // split outputs also own their zero padding columns
  dim3 grid((ncols + BN - 1) / BN, (p.m + BM - 1) / BM);
  gemm_tcgen05_kernel<BN><<<grid, kGemmThreads, smem, st>>>(ma, mw, p);
  return FB200_OK;
}

}  // namespace fb200

using namespace fb200;

extern "C" __attribute__((visibility("default"))) int fb200_gemm(const void* a, int a_rows, int a_parts, const void* w, int w_parts, int m, int n, int k,
                          int taps, const float* bias, int relu, const float* res1, int ld_res1, const float* res2,
                          int ld_res2, float* out_f32, int ld_out, void* out_split, int out_parts, int impl,
                          void* stream) {
  FB_CHECK_ARG(m >= 0 && n > 0 && k > 0, "gemm: bad shape m=%d n=%d k=%d", m, n, k);
  FB_CHECK_ARG((a_parts == 1 || a_parts == 3) && (w_parts == 1 || w_parts == 3), "gemm: parts must be 1 or 3");
  FB_CHECK_ARG(taps >= 1 && taps <= 3, "gemm: taps %d unsupported", taps);
  FB_CHECK_ARG(a_rows >= m + taps - 1 || taps == 1, "gemm: a_rows %d too small for m=%d taps=%d", a_rows, m, taps);
  FB_CHECK_ARG(out_f32 != nullptr || out_split != nullptr, "gemm: no output");
  FB_CHECK_ARG(out_split == nullptr || out_parts == 1 || out_parts == 3, "gemm: out_parts must be 1 or 3");
  if (m == 0) return FB200_OK;
  ProfScope prof(kProfGemm, stream, 2.0 * (double)m * n * k * taps);
  GemmParams p;
  p.m = m;
  p.n = n;
  p.kp_a = p.kp_w = round_up(k, BK);
  p.k_blocks = p.kp_a / BK;
  p.taps = taps;
  if (a_parts == 3 && w_parts == 3) {
    // hi=0, mid=1, lo=2; smallest products first so the fp32 accumulator keeps their bits
    const int ta[6] = {1, 0, 2, 0, 1, 0}, tb[6] = {1, 2, 0, 1, 0, 0};
    p.nterms = 6;
    for (int i = 0; i < 6; ++i) p.term_a[i] = ta[i], p.term_b[i] = tb[i];
  } else {
    FB_CHECK_ARG(a_parts == 1 && w_parts == 1, "gemm: mixed split counts (%d, %d) unsupported", a_parts, w_parts);
    p.nterms = 1;
    for (int i = 0; i < 6; ++i) p.term_a[i] = p.term_b[i] = 0;
  }
  p.bias = bias;
  p.relu = relu;
  p.res1 = res1;
  p.ld_res1 = ld_res1;
  p.res2 = res2;
  p.ld_res2 = ld_res2;
  p.out_f32 = out_f32;
  p.ld_out = ld_out;
  p.out_split = (__nv_bfloat16*)out_split;
  p.out_parts = out_split ? out_parts : 0;
  p.out_kp = out_split ? round_up(n, BK) : 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int lda = a_parts * p.kp_a, ldw = w_parts * p.kp_w;

  if (impl == 1) {
    const int ncols = p.out_split ? p.out_kp : n;
    dim3 grid((ncols + 31) / 32, (m + 127) / 128);
    gemm_simt_kernel<<<grid, 128, 0, st>>>((const __nv_bfloat16*)a, a_rows, lda, (const __nv_bfloat16*)w, ldw, p);
    FB_CHECK_LAUNCH("gemm_simt");
    return FB200_OK;
  }
  FB_CHECK_ARG(impl == 0, "gemm: impl %d unknown", impl);
  CUtensorMap ma, mw;
  int rc = make_map(&ma, a, (uint64_t)a_rows, (uint64_t)lda, BM);
  if (rc) return rc;
  rc = make_map(&mw, w, (uint64_t)taps * n, (uint64_t)ldw, 128);
  if (rc) return rc;
  rc = launch_tcgen05<128>(ma, mw, p, st);
  if (rc) return rc;
  FB_CHECK_LAUNCH("gemm_tcgen05");
  return FB200_OK;
}
