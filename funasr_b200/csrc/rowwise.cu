// Row-wise HBM-bound kernels: LayerNorm (fp32 and/or bf16 split-operand output), operand
// splitting, FSMN depthwise conv with per-sequence zero halos.
#include "common.cuh"
#include "sm100_ptx.cuh"
#include "stack.cuh"

namespace fb200 {

constexpr int kLnWarps = 4;

// One warp per row, the row lives in registers (NV float4 per lane), two-pass mean/variance
// like torch.nn.LayerNorm.  dim % 4 == 0, dim <= NV*128.
template <int NV>
__global__ void __launch_bounds__(kLnWarps * 32)
layernorm_kernel(const float* __restrict__ x, int ld_x, int rows, int dim, const float* __restrict__ gamma,
                 const float* __restrict__ beta, float eps, float* __restrict__ out_f32, int ld_f32,
                 __nv_bfloat16* __restrict__ out_split, int nparts, int halo,
                 const int32_t* __restrict__ cu_rows, int batch) {
  pdl_trigger();
  pdl_wait();  // x comes from the previous kernel
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * kLnWarps + warp;
  if (row >= rows) return;
  const int nvec = dim >> 2;
  const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * ld_x);
  float4 v[NV];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int idx = lane + i * 32;
    v[i] = idx < nvec ? xr[idx] : make_float4(0.f, 0.f, 0.f, 0.f);
    s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
  }
  const float mean = warp_sum(s) / (float)dim;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int idx = lane + i * 32;
    if (idx < nvec) {
      const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
      q += (a * a + b * b) + (c * c + d * d);
    }
  }
  const float rstd = rsqrtf(warp_sum(q) / (float)dim + eps);

  int out_row = row;
  const int kp = round_up(dim, 64);
  if (halo) {
    const int b = find_seq(cu_rows, batch, row);
    out_row = row + 1 + b;
    // the zero rows of the halo layout are owned by the first row of each sequence (the row
    // before it) and by the very last row (the row after it)
    if (out_split) {
      const uint4 z = make_uint4(0u, 0u, 0u, 0u);
      const int nv16 = nparts * kp / 8;
      if (row == __ldg(cu_rows + b)) {
        uint4* zr = reinterpret_cast<uint4*>(out_split + (size_t)(out_row - 1) * (nparts * kp));
        for (int c = lane; c < nv16; c += 32) zr[c] = z;
      }
      if (row == rows - 1) {
        uint4* zr = reinterpret_cast<uint4*>(out_split + (size_t)(out_row + 1) * (nparts * kp));
        for (int c = lane; c < nv16; c += 32) zr[c] = z;
      }
    }
  }
  __nv_bfloat16* srow = out_split ? out_split + (size_t)out_row * (nparts * kp) : nullptr;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int idx = lane + i * 32;
    if (idx < nvec) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + idx);
      const float4 b = __ldg(reinterpret_cast<const float4*>(beta) + idx);
      float4 y;
      y.x = (v[i].x - mean) * rstd * g.x + b.x;
      y.y = (v[i].y - mean) * rstd * g.y + b.y;
      y.z = (v[i].z - mean) * rstd * g.z + b.z;
      y.w = (v[i].w - mean) * rstd * g.w + b.w;
      if (out_f32) reinterpret_cast<float4*>(out_f32 + (size_t)row * ld_f32)[idx] = y;
      if (srow) {
        const float yy[4] = {y.x, y.y, y.z, y.w};
        store_split4(srow, kp, idx * 4, yy, nparts);
      }
    }
  }
  if (srow) {  // zero the K padding
    for (int c = dim + lane; c < kp; c += 32)
      for (int pp = 0; pp < nparts; ++pp) srow[pp * kp + c] = __float2bfloat16_rn(0.f);
  }
}

// fp32 [rows, dim] -> bf16 split [rows, nparts*Kp]; one thread per (row, column of Kp).
// cu_rows != NULL: halo layout (row r lands at r + 1 + seq(r); the interleaved rows are zeroed).
__global__ void split_rows_kernel(const float* __restrict__ x, int ld_x, int rows, int dim,
                                  __nv_bfloat16* __restrict__ out, int nparts, const int32_t* __restrict__ cu_rows,
                                  int batch) {
  pdl_trigger();
  pdl_wait();
  const int kp = round_up(dim, 64);
  const size_t total = (size_t)rows * kp;
  const size_t ld = (size_t)nparts * kp;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / kp), c = (int)(i - (size_t)r * kp);
    const float val = c < dim ? x[(size_t)r * ld_x + c] : 0.f;
    size_t orow = r;
    if (cu_rows != nullptr) {
      const int b = find_seq(cu_rows, batch, r);
      orow = (size_t)r + 1 + b;
      if (r == __ldg(cu_rows + b)) store_split(out + (orow - 1) * ld, kp, c, 0.f, nparts);
      if (r == rows - 1) store_split(out + (orow + 1) * ld, kp, c, 0.f, nparts);
    }
    store_split(out + orow * ld, kp, c, val, nparts);
  }
}

constexpr int kFsmnRows = 64;    // output rows per CTA at most: 18944 rows = 296 CTAs = two per SM, one wave
constexpr int kFsmnMaxK = 32;
constexpr int kFsmnBlock = 8;    // rows per ring stage
constexpr int kFsmnStages = 3;   // ring depth: 3 x (8 input rows + 8 residual rows) x 2 KB = 96 KB per CTA
constexpr int kFsmnCols = 512;   // channels per CTA (128 consumer threads x 4)
constexpr int kFsmnThreads = 160;  // 4 consumer warps + 1 copy warp
constexpr int kFsmnStageBytes = 2 * kFsmnBlock * kFsmnCols * 4;
constexpr int kFsmnSmemBytes = kFsmnStages * kFsmnStageBytes + 1024;

// out[r, c] = sum_j w[c, j] * v[r + j - left, c] (inside the sequence of r) + v[r, c] (+ res[r, c]).
// One CTA = kFsmnRows consecutive packed rows x 512 channels, streamed once: a copy warp moves 8-row blocks of the
// input (and of the residual) global -> shared memory with bulk copies behind mbarriers (up to 96 KB in flight per
// CTA, no registers tied up), the 4 consumer warps walk down the rows with the last KS inputs of their 4 channels
// in registers (sliding window).  Every input element is read once (+ a (KS-1)-row halo per CTA).
// Stream position t: input row r0 - LEFT + t arrives, output row r0 + t - (KS - 1) leaves.
template <int KS>
__global__ void __launch_bounds__(kFsmnThreads)
fsmn_stream_kernel(const float* __restrict__ v, int ld_v, const int32_t* __restrict__ cu_rows, int batch,
                   int total_rows, int dim, const float* __restrict__ weight, const float* __restrict__ res,
                   int ld_res, float* __restrict__ out, int ld_out, int cta_rows) {
  constexpr int LEFT = (KS - 1) / 2;
  constexpr uint32_t FULL = (1u << KS) - 1u;
  pdl_trigger();
  pdl_wait();  // v and the residual come from the previous kernels
  extern __shared__ __align__(128) uint8_t fsmn_smem[];
  __shared__ uint32_t s_mask[kFsmnRows];  // bit j: tap j of this output row lies inside its sequence
  __shared__ __align__(8) uint64_t s_bar[2 * kFsmnStages];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r0 = blockIdx.x * cta_rows, c0 = blockIdx.y * kFsmnCols;
  const int nrows = min(cta_rows, total_rows - r0);
  const int ncols = min(kFsmnCols, dim - c0);
  const uint32_t row_bytes = (uint32_t)ncols * 4u;
  const int nblocks = (nrows + KS - 1 + kFsmnBlock - 1) / kFsmnBlock;
  const uint32_t ring = (smem_u32(fsmn_smem) + 127u) & ~127u;
  auto bar_full = [&](int s) { return smem_u32(&s_bar[s]); };
  auto bar_empty = [&](int s) { return smem_u32(&s_bar[kFsmnStages + s]); };
  if (threadIdx.x < kFsmnRows) {
    uint32_t m = 0;
    if ((int)threadIdx.x < nrows) {
      const int r = r0 + threadIdx.x;
      const int b = find_seq(cu_rows, batch, r);
      const int beg = cu_rows[b], end = cu_rows[b + 1];
#pragma unroll
      for (int j = 0; j < KS; ++j)
        if (r - LEFT + j >= beg && r - LEFT + j < end) m |= 1u << j;
    }
    s_mask[threadIdx.x] = m;
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < kFsmnStages; ++s) {
      mbar_init(bar_full(s), 1);
      mbar_init(bar_empty(s), 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == 4) {
    // ---- copy warp: one lane per row of a block (lanes 0..7 input rows, 8..15 residual rows)
    for (int b = 0; b < nblocks; ++b) {
      const int s = b % kFsmnStages, use = b / kFsmnStages;
      if (use > 0) mbar_wait(bar_empty(s), (uint32_t)(use - 1) & 1u);
      const int t0 = b * kFsmnBlock;
      const int g = lane & 7;
      const bool is_res = lane >= 8;
      bool ok = false;
      const float* src = nullptr;
      if (lane < 8) {
        const int rr = r0 - LEFT + t0 + g;  // input row; rows outside the matrix are never inside a sequence
        ok = rr >= 0 && rr < total_rows && t0 + g < nrows + KS - 1;
        src = v + (size_t)rr * ld_v + c0;
      } else if (lane < 16 && res != nullptr) {
        const int i = t0 + g - (KS - 1);  // output row the residual belongs to
        ok = i >= 0 && i < nrows;
        src = res + (size_t)(r0 + i) * ld_res + c0;
      }
      const uint32_t n_ok = __popc(__ballot_sync(0xffffffffu, ok));
      if (lane == 0) mbar_expect_tx(bar_full(s), n_ok * row_bytes);
      __syncwarp();
      if (ok) {
        const uint32_t dst = ring + (uint32_t)s * kFsmnStageBytes + (uint32_t)((is_res ? kFsmnBlock : 0) + g) * (kFsmnCols * 4);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                     "l"(src), "r"(row_bytes), "r"(bar_full(s))
                     : "memory");
      }
    }
    return;
  }

  // ---- consumer warps: thread = 4 channels
  const int c = threadIdx.x * 4;
  const bool active = c < ncols;
  float4 w[KS];
#pragma unroll
  for (int j = 0; j < KS; ++j)
    w[j] = active ? make_float4(__ldg(weight + (size_t)(c0 + c + 0) * KS + j), __ldg(weight + (size_t)(c0 + c + 1) * KS + j),
                                __ldg(weight + (size_t)(c0 + c + 2) * KS + j), __ldg(weight + (size_t)(c0 + c + 3) * KS + j))
                  : make_float4(0.f, 0.f, 0.f, 0.f);
  float4 win[KS];  // win[j] holds input row (r - LEFT + j) while output row r is produced
#pragma unroll
  for (int j = 0; j < KS; ++j) win[j] = make_float4(0.f, 0.f, 0.f, 0.f);
  const bool has_res = res != nullptr;
  for (int b = 0; b < nblocks; ++b) {
    const int s = b % kFsmnStages, use = b / kFsmnStages;
    mbar_wait(bar_full(s), (uint32_t)use & 1u);
    const uint32_t base = ring + (uint32_t)s * kFsmnStageBytes + (uint32_t)c * 4u;
#pragma unroll
    for (int g = 0; g < kFsmnBlock; ++g) {
      const int i = b * kFsmnBlock + g - (KS - 1);
      uint4 xv, rv = make_uint4(0u, 0u, 0u, 0u);
      ld_shared_v4(base + (uint32_t)g * (kFsmnCols * 4), xv);
      if (has_res) ld_shared_v4(base + (uint32_t)(kFsmnBlock + g) * (kFsmnCols * 4), rv);
#pragma unroll
      for (int j = 0; j < KS - 1; ++j) win[j] = win[j + 1];
      win[KS - 1] = make_float4(__uint_as_float(xv.x), __uint_as_float(xv.y), __uint_as_float(xv.z), __uint_as_float(xv.w));
      if (i >= 0 && i < nrows && active) {
        const uint32_t m = s_mask[i];
        float4 conv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (m == FULL) {
#pragma unroll
          for (int j = 0; j < KS; ++j) {
            conv.x = fmaf(w[j].x, win[j].x, conv.x);
            conv.y = fmaf(w[j].y, win[j].y, conv.y);
            conv.z = fmaf(w[j].z, win[j].z, conv.z);
            conv.w = fmaf(w[j].w, win[j].w, conv.w);
          }
        } else {
#pragma unroll
          for (int j = 0; j < KS; ++j)
            if ((m >> j) & 1u) {
              conv.x = fmaf(w[j].x, win[j].x, conv.x);
              conv.y = fmaf(w[j].y, win[j].y, conv.y);
              conv.z = fmaf(w[j].z, win[j].z, conv.z);
              conv.w = fmaf(w[j].w, win[j].w, conv.w);
            }
        }
        float4 acc = win[LEFT];  // + v (the identity branch)
        if (has_res) {
          acc.x += __uint_as_float(rv.x); acc.y += __uint_as_float(rv.y);
          acc.z += __uint_as_float(rv.z); acc.w += __uint_as_float(rv.w);
        }
        acc.x += conv.x; acc.y += conv.y; acc.z += conv.z; acc.w += conv.w;
        *reinterpret_cast<float4*>(out + (size_t)(r0 + i) * ld_out + c0 + c) = acc;
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_empty(s));
  }
}

// Generic tap count (runtime): one thread = 4 channels of one row, taps re-read through L1.
__global__ void fsmn_generic_kernel(const float* __restrict__ v, int ld_v, const int32_t* __restrict__ cu_rows,
                                    int batch, int total_rows, int dim, const float* __restrict__ weight, int ksize,
                                    const float* __restrict__ res, int ld_res, float* __restrict__ out, int ld_out) {
  __shared__ int s_beg, s_end;
  const int r = blockIdx.x;
  if (threadIdx.x == 0) {
    const int b = find_seq(cu_rows, batch, r);
    s_beg = cu_rows[b];
    s_end = cu_rows[b + 1];
  }
  __syncthreads();
  const int left = (ksize - 1) / 2;
  for (int c = threadIdx.x * 4; c < dim; c += blockDim.x * 4) {
    float4 acc = *reinterpret_cast<const float4*>(v + (size_t)r * ld_v + c);
    if (res) {
      const float4 rr = *reinterpret_cast<const float4*>(res + (size_t)r * ld_res + c);
      acc.x += rr.x; acc.y += rr.y; acc.z += rr.z; acc.w += rr.w;
    }
    for (int j = 0; j < ksize; ++j) {
      const int rr = r + j - left;
      if (rr >= s_beg && rr < s_end) {
        const float4 x = *reinterpret_cast<const float4*>(v + (size_t)rr * ld_v + c);
        acc.x = fmaf(__ldg(weight + (size_t)(c + 0) * ksize + j), x.x, acc.x);
        acc.y = fmaf(__ldg(weight + (size_t)(c + 1) * ksize + j), x.y, acc.y);
        acc.z = fmaf(__ldg(weight + (size_t)(c + 2) * ksize + j), x.z, acc.z);
        acc.w = fmaf(__ldg(weight + (size_t)(c + 3) * ksize + j), x.w, acc.w);
      }
    }
    *reinterpret_cast<float4*>(out + (size_t)r * ld_out + c) = acc;
  }
}

// padded [batch, max_len, dim] <-> packed [sum len, dim]; one CTA per row, float4 per thread.
__global__ void pack_rows_kernel(const float* __restrict__ padded, long long batch_stride, int row_stride,
                                 const int32_t* __restrict__ cu_rows, int batch, int dim,
                                 float* __restrict__ packed, int ld_packed) {
  __shared__ int s_b;
  const int r = blockIdx.x;
  if (threadIdx.x == 0) s_b = find_seq(cu_rows, batch, r);
  __syncthreads();
  const int b = s_b, t = r - cu_rows[b];
  const float* src = padded + (size_t)b * batch_stride + (size_t)t * row_stride;
  for (int c = threadIdx.x * 4; c < dim; c += blockDim.x * 4)
    *reinterpret_cast<float4*>(packed + (size_t)r * ld_packed + c) = *reinterpret_cast<const float4*>(src + c);
}

__global__ void unpack_rows_kernel(const float* __restrict__ packed, int ld_packed, const int32_t* __restrict__ cu_rows,
                                   int max_len, int dim, float pad_value, float* __restrict__ padded) {
  const int b = blockIdx.y, t = blockIdx.x;
  const int beg = cu_rows[b], len = cu_rows[b + 1] - beg;
  float* dst = padded + ((size_t)b * max_len + t) * dim;
  if (t < len) {
    const float* src = packed + (size_t)(beg + t) * ld_packed;
    for (int c = threadIdx.x * 4; c < dim; c += blockDim.x * 4)
      *reinterpret_cast<float4*>(dst + c) = *reinterpret_cast<const float4*>(src + c);
  } else {
    const float4 pv = make_float4(pad_value, pad_value, pad_value, pad_value);
    for (int c = threadIdx.x * 4; c < dim; c += blockDim.x * 4) *reinterpret_cast<float4*>(dst + c) = pv;
  }
}

}  // namespace fb200

using namespace fb200;

extern "C" __attribute__((visibility("default"))) int fb200_pack_rows(const float* padded, long long batch_stride,
                                                                       int row_stride, const int32_t* cu_rows, int batch,
                                                                       int total_rows, int dim, float* packed,
                                                                       int ld_packed, void* stream) {
  FB_CHECK_ARG(dim % 4 == 0 && row_stride % 4 == 0 && batch_stride % 4 == 0 && ld_packed % 4 == 0,
               "pack_rows: dim and strides must be multiples of 4");
  if (total_rows == 0) return FB200_OK;
  ProfScope prof(kProfSplit, stream, 8.0 * (double)total_rows * dim);
  pack_rows_kernel<<<total_rows, 128, 0, (cudaStream_t)stream>>>(padded, batch_stride, row_stride, cu_rows, batch, dim,
                                                                 packed, ld_packed);
  FB_CHECK_LAUNCH("pack_rows");
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_unpack_rows(const float* packed, int ld_packed,
                                                                         const int32_t* cu_rows, int batch, int max_len,
                                                                         int dim, float pad_value, float* padded,
                                                                         void* stream) {
  FB_CHECK_ARG(dim % 4 == 0 && ld_packed % 4 == 0, "unpack_rows: dim and ld must be multiples of 4");
  FB_CHECK_ARG(batch > 0 && batch <= 65535, "unpack_rows: batch %d out of range", batch);
  if (max_len == 0) return FB200_OK;
  ProfScope prof(kProfSplit, stream, 8.0 * (double)batch * max_len * dim);
  unpack_rows_kernel<<<dim3(max_len, batch), 128, 0, (cudaStream_t)stream>>>(packed, ld_packed, cu_rows, max_len, dim,
                                                                             pad_value, padded);
  FB_CHECK_LAUNCH("unpack_rows");
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_layernorm(const float* x, int ld_x, int rows, int dim, const float* gamma, const float* beta,
                               float eps, float* out_f32, int ld_f32, void* out_split, int nparts, int halo,
                               const int32_t* cu_rows, int batch, void* stream) {
  FB_CHECK_ARG(dim % 4 == 0 && dim <= 2048 && ld_x % 4 == 0, "layernorm: dim %d / ld %d unsupported", dim, ld_x);
  FB_CHECK_ARG(out_f32 == nullptr || ld_f32 % 4 == 0, "layernorm: ld_f32 must be a multiple of 4");
  FB_CHECK_ARG(nparts >= 1 && nparts <= 3, "layernorm: nparts must be 1, 2 or 3");
  FB_CHECK_ARG(!halo || (cu_rows != nullptr && batch > 0), "layernorm: halo output needs cu_rows");
  if (rows == 0) return FB200_OK;
  if (stack_recording())
    return stack_record_layernorm(x, ld_x, rows, dim, gamma, beta, eps, out_f32, ld_f32, out_split, nparts, halo, cu_rows, batch);
  ProfScope prof(kProfLayerNorm, stream, (double)rows * dim * (4.0 + (out_f32 ? 4.0 : 0.0) + (out_split ? 2.0 * nparts : 0.0)));
  const int grid = (rows + kLnWarps - 1) / kLnWarps;
  cudaStream_t st = (cudaStream_t)stream;
  __nv_bfloat16* sp = (__nv_bfloat16*)out_split;
#define LN_LAUNCH(NV)                                                                                            \
  launch_pdl(layernorm_kernel<NV>, dim3(grid), dim3(kLnWarps * 32), 0, st, 1, x, ld_x, rows, dim, gamma, beta, eps, out_f32, \
             ld_f32, sp, nparts, halo, cu_rows, batch)
  if (dim <= 512) LN_LAUNCH(4);
  else if (dim <= 640) LN_LAUNCH(5);
  else LN_LAUNCH(16);
#undef LN_LAUNCH
  FB_CHECK_LAUNCH("layernorm");
  return FB200_OK;
}

static int split_rows_impl(const float* x, int ld_x, int rows, int dim, void* out_split, int nparts,
                           const int32_t* cu_rows, int batch, void* stream) {
  FB_CHECK_ARG(nparts >= 1 && nparts <= 3, "split_rows: nparts must be 1, 2 or 3");
  if (rows == 0) return FB200_OK;
  ProfScope prof(kProfSplit, stream, (double)rows * dim * (4.0 + 2.0 * nparts));
  const size_t total = (size_t)rows * round_up(dim, 64);
  int grid = (int)((total + 255) / 256);
  if (grid > 148 * 16) grid = 148 * 16;
  launch_pdl(split_rows_kernel, dim3(grid), dim3(256), 0, stream, 1, x, ld_x, rows, dim, (__nv_bfloat16*)out_split, nparts,
             cu_rows, batch);
  FB_CHECK_LAUNCH("split_rows");
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_split_rows(const float* x, int ld_x, int rows, int dim,
                                                                        void* out_split, int nparts, void* stream) {
  return split_rows_impl(x, ld_x, rows, dim, out_split, nparts, nullptr, 0, stream);
}

extern "C" __attribute__((visibility("default"))) int fb200_split_rows_halo(const float* x, int ld_x, int rows, int dim,
                                                                             void* out_split, int nparts,
                                                                             const int32_t* cu_rows, int batch,
                                                                             void* stream) {
  FB_CHECK_ARG(cu_rows != nullptr && batch > 0, "split_rows_halo: cu_rows / batch missing");
  return split_rows_impl(x, ld_x, rows, dim, out_split, nparts, cu_rows, batch, stream);
}

extern "C" __attribute__((visibility("default"))) int fb200_split_weight(const float* w, int n, int k, void* out_split, int nparts, void* stream) {
  return fb200_split_rows(w, k, n, k, out_split, nparts, stream);
}

extern "C" __attribute__((visibility("default"))) int fb200_fsmn(const float* v, int ld_v, const int32_t* cu_rows, int batch, int total_rows, int dim,
                          const float* weight, int ksize, const float* residual, int ld_res, float* out, int ld_out,
                          void* stream) {
  FB_CHECK_ARG(dim % 4 == 0 && ld_v % 4 == 0 && ld_out % 4 == 0, "fsmn: dim/ld must be multiples of 4");
  FB_CHECK_ARG(ksize >= 1 && ksize <= kFsmnMaxK, "fsmn: ksize %d unsupported", ksize);
  FB_CHECK_ARG(residual == nullptr || ld_res % 4 == 0, "fsmn: ld_res must be a multiple of 4");
  if (total_rows == 0) return FB200_OK;
  if (stack_recording())
    return stack_record_fsmn(v, ld_v, cu_rows, batch, total_rows, dim, weight, ksize, residual, ld_res, out, ld_out);
  ProfScope prof(kProfFsmn, stream, (double)total_rows * dim * 4.0 * (residual ? 3.0 : 2.0));
  if (ksize == 11 || ksize == 21) {  // 11: Paraformer / SenseVoice; 21: the SeACo decoder (seaco_paraformer/template.yaml)
    // rows per CTA: 64 when that fills the GPU (two CTAs per SM), fewer for small batches (a CTA walks its rows
    // serially, so short walks are what keeps a small launch short; the halo re-read grows to (rows + K - 1) / rows)
    int cta_rows = (total_rows + 2 * 148 - 1) / (2 * 148);
    cta_rows = (cta_rows + kFsmnBlock - 1) / kFsmnBlock * kFsmnBlock;
    cta_rows = cta_rows < 16 ? 16 : (cta_rows > kFsmnRows ? kFsmnRows : cta_rows);
    const dim3 grid((total_rows + cta_rows - 1) / cta_rows, (dim + kFsmnCols - 1) / kFsmnCols);
#define FSMN_STREAM(KS)                                                                                                      \
  do {                                                                                                                       \
    static bool attr_set = false; /* per process; idempotent */                                                              \
    if (!attr_set) {                                                                                                         \
      cudaError_t e = cudaFuncSetAttribute(fsmn_stream_kernel<KS>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFsmnSmemBytes); \
      if (e != cudaSuccess) {                                                                                                \
        set_error("fsmn: cudaFuncSetAttribute: %s", cudaGetErrorString(e));                                                  \
        return FB200_ECUDA;                                                                                                  \
      }                                                                                                                      \
      attr_set = true;                                                                                                       \
    }                                                                                                                        \
    launch_pdl(fsmn_stream_kernel<KS>, grid, dim3(kFsmnThreads), kFsmnSmemBytes, stream, 1,                                 \
               v, ld_v, cu_rows, batch, total_rows, dim, weight, residual, ld_res, out, ld_out, cta_rows);                  \
  } while (0)
    if (ksize == 11) FSMN_STREAM(11);
    else FSMN_STREAM(21);
#undef FSMN_STREAM
  }
  else
    fsmn_generic_kernel<<<total_rows, 128, 0, (cudaStream_t)stream>>>(v, ld_v, cu_rows, batch, total_rows, dim, weight,
                                                                      ksize, residual, ld_res, out, ld_out);
  FB_CHECK_LAUNCH("fsmn");
  return FB200_OK;
}
