// Kernels of the Conformer encoder path (BASELINE config 4) that are not dense contractions:
//
//   conv2d_sub1_kernel      first 3x3 stride-2 convolution of Conv2dSubsampling (1 -> C channels, K = 9: CUDA cores),
//                           ReLU, written as a bf16 split operand in FOUR PARITY PLANES so that the second
//                           3x3 stride-2 convolution becomes a 9-tap tcgen05 GEMM with per-tap row offsets
//                           (no im2col matrix ever exists)
//   gather_rows_kernel      drops the one scratch row per utterance the plane layout carries
//   glu_dwconv_kernel       ConvolutionModule between its two pointwise GEMMs: GLU -> depthwise conv k (zero halo per
//                           utterance) -> BatchNorm (eval, folded to scale/shift) -> swish -> bf16 split operand
//   relpos_attention_kernel legacy relative-position attention (scores = (q+u).k + rel_shift((q+v).p)), streaming
//                           softmax; the (q+v).p products come from a tcgen05 GEMM, the shift is pure indexing
//
// Plane layout (per utterance b: F_b fbank frames, T1 = (F_b-3)/2+1, T2 = (T1-3)/2+1, segment of PT_b = T2+1 rows;
// F1 = (n_mels-3)/2+1 = 39 frequencies after conv 1, FP = (F1+1)/2 = 20 slots per segment row):
//   plane (pt, pf), row (seg_b + t') * FP + f'  holds conv1 output (t1 = 2t'+pt, f1 = 2f'+pf), zero where t1 >= T1 or
//   f1 >= F1.  conv2 output (t2, f2) = sum_{dt,df} W[dt,df] . plane(dt&1, df&1)[row(t2,f2) + (dt>>1)*FP + (df>>1)].
#include "common.cuh"

namespace fb200 {

// ---------------------------------------------------------------- conv 1 -> parity planes ----
constexpr int kSubThreads = 128;  // 2 channels per thread: C <= 256

__global__ void __launch_bounds__(kSubThreads)
conv2d_sub1_kernel(const float* __restrict__ feats, int ld_f, const int32_t* __restrict__ cu_frames,
                   const int32_t* __restrict__ cu_seg, int batch, int n_mels, const float* __restrict__ weight,
                   const float* __restrict__ bias, int C, __nv_bfloat16* __restrict__ planes, int nparts, long long plane_rows,
                   int FP, int F1) {
  extern __shared__ float xs[];  // [5][n_mels]
  const int seg = blockIdx.x;
  const int b = find_seq(cu_seg, batch, seg);
  const int tp = seg - __ldg(cu_seg + b);
  const int f_beg = __ldg(cu_frames + b), F = __ldg(cu_frames + b + 1) - f_beg;
  const int T1 = F >= 3 ? (F - 3) / 2 + 1 : 0;
  for (int i = threadIdx.x; i < 5 * n_mels; i += kSubThreads) {
    const int r = 4 * tp + i / n_mels;
    xs[i] = r < F ? __ldg(feats + (size_t)(f_beg + r) * ld_f + i % n_mels) : 0.f;
  }
  __syncthreads();
  const int c0 = 2 * threadIdx.x;
  if (c0 >= C) return;
  float w0[9], w1[9];
#pragma unroll
  for (int j = 0; j < 9; ++j) {
    w0[j] = __ldg(weight + (size_t)c0 * 9 + j);
    w1[j] = __ldg(weight + (size_t)(c0 + 1) * 9 + j);
  }
  const float b0 = __ldg(bias + c0), b1 = __ldg(bias + c0 + 1);
  const int ld = nparts * C;
  for (int pl = 0; pl < 4; ++pl) {
    const int pt = pl >> 1, pf = pl & 1;
    const bool t_ok = 2 * tp + pt < T1;
    __nv_bfloat16* prow = planes + ((size_t)pl * plane_rows + (size_t)seg * FP) * ld + c0;
    for (int fp = 0; fp < FP; ++fp, prow += ld) {
      const int f1 = 2 * fp + pf;
      float y0 = 0.f, y1 = 0.f;
      if (t_ok && f1 < F1) {
        float a0 = 0.f, a1 = 0.f;
#pragma unroll
        for (int dt = 0; dt < 3; ++dt)
#pragma unroll
          for (int df = 0; df < 3; ++df) {
            const float x = xs[(2 * pt + dt) * n_mels + 2 * f1 + df];
            a0 = fmaf(w0[dt * 3 + df], x, a0);
            a1 = fmaf(w1[dt * 3 + df], x, a1);
          }
        y0 = fmaxf(a0 + b0, 0.f);
        y1 = fmaxf(a1 + b1, 0.f);
      }
      const __nv_bfloat162 h = __floats2bfloat162_rn(y0, y1);
      *reinterpret_cast<__nv_bfloat162*>(prow) = h;
      if (nparts > 1) {
        const float r0 = y0 - __low2float(h), r1 = y1 - __high2float(h);
        const __nv_bfloat162 m = __floats2bfloat162_rn(r0, r1);
        *reinterpret_cast<__nv_bfloat162*>(prow + C) = m;
        if (nparts > 2)
          *reinterpret_cast<__nv_bfloat162*>(prow + 2 * C) = __floats2bfloat162_rn(r0 - __low2float(m), r1 - __high2float(m));
      }
    }
  }
}

// ---------------------------------------------------------------- row gather -----------------
__global__ void gather_rows_kernel(const float* __restrict__ src, int ld_src, const int32_t* __restrict__ cu_src,
                                   float* __restrict__ dst, int ld_dst, const int32_t* __restrict__ cu_dst, int batch,
                                   int rows, int dim) {
  const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int lane = threadIdx.x & 31;
  const int b = find_seq(cu_dst, batch, r);
  const size_t s = (size_t)(__ldg(cu_src + b) + r - __ldg(cu_dst + b));
  const float4* sp = reinterpret_cast<const float4*>(src + s * ld_src);
  float4* dp = reinterpret_cast<float4*>(dst + (size_t)r * ld_dst);
  for (int i = lane; i < dim / 4; i += 32) dp[i] = __ldg(sp + i);
}

// ---------------------------------------------------------------- GLU + depthwise + BN + swish
// One CTA = kDwRows consecutive PACKED rows x all channels (the grid is rows / kDwRows whatever the utterance
// lengths: 18 944 rows = one full wave of 2 CTAs per SM).  Phase 1: GLU of the kDwRows + K - 1 input rows into shared
// memory, every thread busy; phase 2: thread = (kDwRpt rows, channel pair), weights and input window in registers,
// taps that fall outside the row's own utterance masked (= the zero padding of the reference's Conv1d), BatchNorm
// (folded), swish, bf16 split store.  (A first version walked the rows serially from global memory with a register
// window per thread: 41 us per 18 944-row call at 9 % of the warp slots -- latency-bound; a second read the strided
// weights through __ldg per tap: 85 us, L1-bound; tiling per utterance left a nearly empty second wave: 47 us.)
constexpr int kDwRows = 64;
constexpr int kDwRpt = 16;
constexpr int kDwThreads = 512;

template <int K>
__global__ void __launch_bounds__(kDwThreads)
glu_dwconv_kernel(const float* __restrict__ g, int ld_g, const int32_t* __restrict__ cu_rows, int batch, int total_rows, int C,
                  const float* __restrict__ weight, const float* __restrict__ bias, const float* __restrict__ bn_scale,
                  const float* __restrict__ bn_shift, __nv_bfloat16* __restrict__ out_split, int nparts, int out_kp) {
  constexpr int PAD = (K - 1) / 2;
  extern __shared__ float glu[];  // [kDwRows + K - 1][C]
  __shared__ int tap_lo[kDwRows], tap_hi[kDwRows];  // taps [lo, hi) of a row stay inside its utterance
  const int g0 = blockIdx.x * kDwRows;
  const int rows = min(kDwRows, total_rows - g0);
  if (threadIdx.x < rows) {
    const int gr = g0 + threadIdx.x;
    int lo = 0, hi = batch;  // largest b with cu_rows[b] <= gr
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (__ldg(cu_rows + mid) <= gr) lo = mid; else hi = mid;
    }
    const int beg = __ldg(cu_rows + lo), end = __ldg(cu_rows + lo + 1);
    tap_lo[threadIdx.x] = max(0, PAD - (gr - beg));
    tap_hi[threadIdx.x] = min(K, PAD + (end - gr));
  }
  const int half = C >> 1;
  const int items = (kDwRows + K - 1) * half;
#pragma unroll 4
  for (int i = threadIdx.x; i < items; i += kDwThreads) {
    const int rr = i / half, c0 = (i - rr * half) * 2;
    const int t = g0 - PAD + rr;
    const bool ok = t >= 0 && t < total_rows;
    const float* gp = g + (size_t)(ok ? t : 0) * ld_g + c0;
    const float2 a = __ldg(reinterpret_cast<const float2*>(gp));
    const float2 gate = __ldg(reinterpret_cast<const float2*>(gp + C));
    float2 v;
    v.x = ok ? __fdividef(a.x, 1.f + __expf(-gate.x)) : 0.f;
    v.y = ok ? __fdividef(a.y, 1.f + __expf(-gate.y)) : 0.f;
    *reinterpret_cast<float2*>(glu + rr * C + c0) = v;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < (kDwRows / kDwRpt) * half; i += kDwThreads) {
    const int grp = i / half, c0 = (i - grp * half) * 2;
    const int r0 = grp * kDwRpt;
    if (r0 >= rows) continue;
    float w0[K], w1[K];
#pragma unroll
    for (int j = 0; j < K; ++j) {
      w0[j] = __ldg(weight + (size_t)c0 * K + j);
      w1[j] = __ldg(weight + (size_t)(c0 + 1) * K + j);
    }
    const float2 bi = *reinterpret_cast<const float2*>(bias + c0);
    const float2 sc = *reinterpret_cast<const float2*>(bn_scale + c0);
    const float2 sh = *reinterpret_cast<const float2*>(bn_shift + c0);
    constexpr bool kWindow = K <= 15;
    float2 win[kWindow ? kDwRpt + K - 1 : 1];
    if constexpr (kWindow) {
#pragma unroll
      for (int j = 0; j < kDwRpt + K - 1; ++j) win[j] = *reinterpret_cast<const float2*>(glu + (r0 + j) * C + c0);
    }
#pragma unroll
    for (int rr = 0; rr < kDwRpt; ++rr) {
      const int r = r0 + rr;
      if (r >= rows) break;
      const int lo = tap_lo[r], hi = tap_hi[r];
      float y0 = 0.f, y1 = 0.f;
      if (lo == 0 && hi == K) {  // interior row (the whole warp shares r): no masking
#pragma unroll
        for (int j = 0; j < K; ++j) {
          const float2 x = kWindow ? win[kWindow ? rr + j : 0] : *reinterpret_cast<const float2*>(glu + (r + j) * C + c0);
          y0 = fmaf(w0[j], x.x, y0);
          y1 = fmaf(w1[j], x.y, y1);
        }
      } else {
#pragma unroll
        for (int j = 0; j < K; ++j) {
          float2 x = kWindow ? win[kWindow ? rr + j : 0] : *reinterpret_cast<const float2*>(glu + (r + j) * C + c0);
          if (j < lo || j >= hi) x = make_float2(0.f, 0.f);
          y0 = fmaf(w0[j], x.x, y0);
          y1 = fmaf(w1[j], x.y, y1);
        }
      }
      y0 = (y0 + bi.x) * sc.x + sh.x;
      y1 = (y1 + bi.y) * sc.y + sh.y;
      y0 = __fdividef(y0, 1.f + __expf(-y0));
      y1 = __fdividef(y1, 1.f + __expf(-y1));
      __nv_bfloat16* orow = out_split + (size_t)(g0 + r) * (nparts * out_kp) + c0;
      const __nv_bfloat162 h = __floats2bfloat162_rn(y0, y1);
      *reinterpret_cast<__nv_bfloat162*>(orow) = h;
      if (nparts > 1) {
        const float e0 = y0 - __low2float(h), e1 = y1 - __high2float(h);
        const __nv_bfloat162 m = __floats2bfloat162_rn(e0, e1);
        *reinterpret_cast<__nv_bfloat162*>(orow + out_kp) = m;
        if (nparts > 2)
          *reinterpret_cast<__nv_bfloat162*>(orow + 2 * out_kp) = __floats2bfloat162_rn(e0 - __low2float(m), e1 - __high2float(m));
      }
    }
  }
}

// ---------------------------------------------------------------- relative-position attention
// fp32 on the CUDA cores over the split operands (rebuilt as hi + mid while staging): one CTA = 64 query rows of
// one (utterance, head), keys / values streamed in tiles of 64, online softmax.  Scores:
//   s[i, j] = ((q_i + u) . k_j + shift(bd)[i, j]) * scale,  bd[i, c] = (q_i + v) . p_c  (from the GEMM),
//   shift(bd)[i, j] = bd[i, T-1-i+j]  (j <= i),  0  (j = i+1),  bd[i+1, j-i-2]  (j >= i+2)
// -- LegacyRelPositionMultiHeadedAttention.rel_shift's pad / view / slice in closed form.
constexpr int RQ = 64, RK = 64, RD = 64;
constexpr int RDP = RD + 4, RPP = RK + 4;
constexpr int kRelThreads = 256;
constexpr size_t kRelSmem = (size_t)(RQ * RDP + RK * RDP + RK * RD + RQ * RPP) * sizeof(float);

__device__ __forceinline__ float4 rel_load_split4(const __nv_bfloat16* p, int part_stride, int parts) {
  float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int pp = 0; pp < parts; ++pp) {
    const uint2 raw = *reinterpret_cast<const uint2*>(p + (size_t)pp * part_stride);
    const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&raw.x);
    const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(&raw.y);
    r.x += __bfloat162float(a.x); r.y += __bfloat162float(a.y);
    r.z += __bfloat162float(b.x); r.w += __bfloat162float(b.y);
  }
  return r;
}

__global__ void __launch_bounds__(kRelThreads)
relpos_attention_kernel(const __nv_bfloat16* __restrict__ q, int ld_q, int q_ps, const __nv_bfloat16* __restrict__ k, int ld_k,
                        const __nv_bfloat16* __restrict__ v, int ld_v, int kv_ps, int in_parts,
                        const float* __restrict__ bd, int ld_bd, long long bd_head_stride,
                        const int32_t* __restrict__ cu_rows, float scale, float* __restrict__ out_f32, int ld_out,
                        __nv_bfloat16* __restrict__ out_split, int nparts, int out_kp) {
  extern __shared__ float smem[];
  float* Qs = smem;               // [RQ][RDP]
  float* Ks = Qs + RQ * RDP;      // [RK][RDP]
  float* Vs = Ks + RK * RDP;      // [RK][RD]
  float* Ps = Vs + RK * RD;       // [RQ][RPP]

  const int b = blockIdx.z, head = blockIdx.y;
  const int beg = cu_rows[b], T = cu_rows[b + 1] - beg;
  const int q0 = blockIdx.x * RQ;
  if (q0 >= T) return;
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int hoff = head * RD;
  const float* bdh = bd + (size_t)head * bd_head_stride;

  for (int i = tid; i < RQ * (RD / 4); i += kRelThreads) {
    const int r = i / (RD / 4), c4 = i % (RD / 4);
    float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
    if (q0 + r < T) val = rel_load_split4(q + (size_t)(beg + q0 + r) * ld_q + hoff + c4 * 4, q_ps, in_parts);
    *reinterpret_cast<float4*>(Qs + r * RDP + c4 * 4) = val;
  }

  float o[4][4];
  float m_run[4], l_run[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    m_run[i] = -INFINITY;
    l_run[i] = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) o[i][j] = 0.f;
  }

  for (int k0 = 0; k0 < T; k0 += RK) {
    __syncthreads();
    for (int i = tid; i < RK * (RD / 4); i += kRelThreads) {
      const int r = i / (RD / 4), c4 = i % (RD / 4);
      float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
      if (k0 + r < T) {
        kv = rel_load_split4(k + (size_t)(beg + k0 + r) * ld_k + hoff + c4 * 4, kv_ps, in_parts);
        vv = rel_load_split4(v + (size_t)(beg + k0 + r) * ld_v + hoff + c4 * 4, kv_ps, in_parts);
      }
      *reinterpret_cast<float4*>(Ks + r * RDP + c4 * 4) = kv;
      *reinterpret_cast<float4*>(Vs + r * RD + c4 * 4) = vv;
    }
    __syncthreads();

    float s[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) s[i][j] = 0.f;
#pragma unroll 4
    for (int d = 0; d < RD; d += 4) {
      float4 qv[4], kv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) qv[i] = *reinterpret_cast<const float4*>(Qs + (ty * 4 + i) * RDP + d);
#pragma unroll
      for (int j = 0; j < 4; ++j) kv[j] = *reinterpret_cast<const float4*>(Ks + (tx + 16 * j) * RDP + d);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          s[i][j] = fmaf(qv[i].x, kv[j].x, s[i][j]);
          s[i][j] = fmaf(qv[i].y, kv[j].y, s[i][j]);
          s[i][j] = fmaf(qv[i].z, kv[j].z, s[i][j]);
          s[i][j] = fmaf(qv[i].w, kv[j].w, s[i][j]);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int qi = q0 + ty * 4 + i;  // query index inside the utterance
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int kj = k0 + tx + 16 * j;
        if (kj >= T || qi >= T) {
          s[i][j] = -INFINITY;
        } else {
          float pos = 0.f;
          if (kj <= qi) pos = __ldg(bdh + (size_t)(beg + qi) * ld_bd + (T - 1 - qi + kj));
          else if (kj >= qi + 2) pos = __ldg(bdh + (size_t)(beg + qi + 1) * ld_bd + (kj - qi - 2));
          s[i][j] = (s[i][j] + pos) * scale;
        }
        mx = fmaxf(mx, s[i][j]);
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
      const float m_new = fmaxf(m_run[i], mx);
      const bool dead = m_new == -INFINITY;  // rows past the utterance end
      const float corr = dead ? 1.f : expf(m_run[i] - m_new);
      float sum = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float pv = dead ? 0.f : expf(s[i][j] - m_new);
        sum += pv;
        Ps[(ty * 4 + i) * RPP + tx + 16 * j] = pv;
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
      l_run[i] = l_run[i] * corr + sum;
      m_run[i] = m_new;
#pragma unroll
      for (int j = 0; j < 4; ++j) o[i][j] *= corr;
    }
    __syncthreads();
#pragma unroll 2
    for (int kk = 0; kk < RK; kk += 4) {
      float4 pv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) pv[i] = *reinterpret_cast<const float4*>(Ps + (ty * 4 + i) * RPP + kk);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float4 v0 = *reinterpret_cast<const float4*>(Vs + (kk + u) * RD + tx * 4);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float pu = u == 0 ? pv[i].x : u == 1 ? pv[i].y : u == 2 ? pv[i].z : pv[i].w;
          o[i][0] = fmaf(pu, v0.x, o[i][0]);
          o[i][1] = fmaf(pu, v0.y, o[i][1]);
          o[i][2] = fmaf(pu, v0.z, o[i][2]);
          o[i][3] = fmaf(pu, v0.w, o[i][3]);
        }
      }
    }
  }

#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = q0 + ty * 4 + i;
    if (r >= T) continue;
    const float inv = 1.f / l_run[i];
    const size_t row = (size_t)(beg + r);
    const int col = hoff + tx * 4;
    const float y[4] = {o[i][0] * inv, o[i][1] * inv, o[i][2] * inv, o[i][3] * inv};
    if (out_f32) *reinterpret_cast<float4*>(out_f32 + row * ld_out + col) = make_float4(y[0], y[1], y[2], y[3]);
    if (out_split) store_split4(out_split + row * (size_t)(nparts * out_kp), out_kp, col, y, nparts);
  }
}

}  // namespace fb200

using namespace fb200;

extern "C" __attribute__((visibility("default"))) int fb200_conv2d_sub1(
    const float* feats, int ld_f, const int32_t* cu_frames, const int32_t* cu_seg, int batch, int total_seg, int n_mels,
    const float* weight, const float* bias, int channels, void* planes, int nparts, void* stream) {
  FB_CHECK_ARG(batch > 0 && total_seg >= 0 && n_mels >= 7, "conv2d_sub1: batch %d / segments %d / n_mels %d", batch, total_seg, n_mels);
  FB_CHECK_ARG(channels % 2 == 0 && channels <= 2 * kSubThreads, "conv2d_sub1: channels %d unsupported (even, <= %d)", channels,
               2 * kSubThreads);
  FB_CHECK_ARG(nparts >= 1 && nparts <= 3, "conv2d_sub1: nparts must be 1, 2 or 3");
  if (total_seg == 0) return FB200_OK;
  const int F1 = (n_mels - 3) / 2 + 1, FP = (F1 + 1) / 2;
  const long long plane_rows = (long long)total_seg * FP;
  ProfScope prof(kProfConvSub, stream, 4.0 * plane_rows * channels * 2.0 * nparts);
  conv2d_sub1_kernel<<<total_seg, kSubThreads, 5 * n_mels * sizeof(float), (cudaStream_t)stream>>>(
      feats, ld_f, cu_frames, cu_seg, batch, n_mels, weight, bias, channels, (__nv_bfloat16*)planes, nparts, plane_rows, FP, F1);
  FB_CHECK_LAUNCH("conv2d_sub1");
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_gather_rows(const float* src, int ld_src, const int32_t* cu_src,
                                                                         float* dst, int ld_dst, const int32_t* cu_dst,
                                                                         int batch, int rows, int dim, void* stream) {
  FB_CHECK_ARG(batch > 0 && rows >= 0 && dim % 4 == 0 && ld_src % 4 == 0 && ld_dst % 4 == 0, "gather_rows: rows %d / dim %d / ld %d %d",
               rows, dim, ld_src, ld_dst);
  if (rows == 0) return FB200_OK;
  ProfScope prof(kProfSplit, stream, 8.0 * rows * dim);
  gather_rows_kernel<<<(rows + 7) / 8, 256, 0, (cudaStream_t)stream>>>(src, ld_src, cu_src, dst, ld_dst, cu_dst, batch, rows, dim);
  FB_CHECK_LAUNCH("gather_rows");
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_glu_dwconv(
    const float* g, int ld_g, const int32_t* cu_rows, int batch, int total_rows, int max_len, int channels, const float* weight,
    int ksize, const float* bias, const float* bn_scale, const float* bn_shift, void* out_split, int nparts, void* stream) {
  FB_CHECK_ARG(batch > 0 && batch <= 65535 && max_len >= 0, "glu_dwconv: batch %d / max_len %d", batch, max_len);
  FB_CHECK_ARG(channels % 2 == 0 && channels <= 1024 && ld_g % 2 == 0 && ld_g >= 2 * channels, "glu_dwconv: channels %d / ld %d",
               channels, ld_g);
  FB_CHECK_ARG(nparts >= 1 && nparts <= 3, "glu_dwconv: nparts must be 1, 2 or 3");
  if (total_rows == 0 || max_len == 0) return FB200_OK;
  ProfScope prof(kProfGluDw, stream, (double)total_rows * channels * (8.0 + 2.0 * nparts));
  const int grid = (total_rows + kDwRows - 1) / kDwRows;
  const int kp = round_up(channels, 64);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = (size_t)(kDwRows + ksize - 1) * channels * sizeof(float);
  FB_CHECK_ARG(smem <= 200 * 1024, "glu_dwconv: %d channels x kernel %d needs %zu B of shared memory", channels, ksize, smem);
#define DW_LAUNCH(K)                                                                                                   \
  do {                                                                                                                 \
    static bool attr_set = false;                                                                                      \
    if (!attr_set && smem > 48 * 1024) {                                                                               \
      cudaError_t e = cudaFuncSetAttribute(glu_dwconv_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); \
      if (e != cudaSuccess) {                                                                                          \
        set_error("glu_dwconv: cudaFuncSetAttribute: %s", cudaGetErrorString(e));                                      \
        return FB200_ECUDA;                                                                                            \
      }                                                                                                                \
      attr_set = true;                                                                                                 \
    }                                                                                                                  \
    glu_dwconv_kernel<K><<<grid, kDwThreads, smem, st>>>(g, ld_g, cu_rows, batch, total_rows, channels, weight, bias,   \
                                                         bn_scale, bn_shift, (__nv_bfloat16*)out_split, nparts, kp);   \
  } while (0)
  if (ksize == 15) DW_LAUNCH(15);
  else if (ksize == 31) DW_LAUNCH(31);
  else if (ksize == 7) DW_LAUNCH(7);
  else {
    set_error("glu_dwconv: kernel size %d unsupported (7, 15, 31)", ksize);
    return FB200_EUNSUPPORTED;
  }
#undef DW_LAUNCH
  FB_CHECK_LAUNCH("glu_dwconv");
  return FB200_OK;
}

namespace fb200 {
// SIMT check kernel of fb200_attention_relpos (impl = 1, and the 3-part mode); entry in relpos_attention_sm100.cu
int relpos_attention_simt(
    const void* q, int ld_q, int q_col0, int q_part_stride, const void* k, int ld_k, int k_col0, const void* v, int ld_v,
    int v_col0, int kv_part_stride, int nparts, const float* bd, int ld_bd, long long bd_head_stride, const int32_t* cu_rows,
    int batch, int rows, int max_len, int heads, int head_dim, float scale, float* out_f32, int ld_out, void* out_split,
    int out_parts, void* stream) {
  FB_CHECK_ARG(head_dim == RD, "attention_relpos: head_dim %d unsupported (64 only)", head_dim);
  FB_CHECK_ARG(nparts >= 1 && nparts <= 3, "attention_relpos: nparts must be 1, 2 or 3");
  FB_CHECK_ARG(ld_q % 4 == 0 && ld_k % 4 == 0 && ld_v % 4 == 0 && q_col0 % 4 == 0 && k_col0 % 4 == 0 && v_col0 % 4 == 0 &&
                   q_part_stride % 4 == 0 && kv_part_stride % 4 == 0,
               "attention_relpos: leading dims / offsets must be multiples of 4");
  FB_CHECK_ARG(out_f32 != nullptr || out_split != nullptr, "attention_relpos: no output");
  FB_CHECK_ARG(out_f32 == nullptr || ld_out % 4 == 0, "attention_relpos: ld_out must be a multiple of 4");
  FB_CHECK_ARG(batch > 0 && batch <= 65535 && heads > 0 && ld_bd >= max_len, "attention_relpos: batch %d / heads %d / ld_bd %d < %d",
               batch, heads, ld_bd, max_len);
  if (rows == 0 || max_len == 0) return FB200_OK;
  double flops = 0.0;  // the caller's max_len bounds it; the exact sum needs the lengths: 4 T^2 d per (utt, head) ~ rows * max_len
  flops = 4.0 * (double)rows * max_len * head_dim * heads;
  ProfScope prof(kProfRelAttention, stream, flops);
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(relpos_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRelSmem);
    if (e != cudaSuccess) {
      set_error("attention_relpos: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return FB200_ECUDA;
    }
    attr_set = true;
  }
  const int out_kp = round_up(heads * head_dim, 64);
  dim3 grid((max_len + RQ - 1) / RQ, heads, batch);
  relpos_attention_kernel<<<grid, kRelThreads, kRelSmem, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)q + q_col0, ld_q, q_part_stride, (const __nv_bfloat16*)k + k_col0, ld_k,
      (const __nv_bfloat16*)v + v_col0, ld_v, kv_part_stride, nparts, bd, ld_bd, bd_head_stride, cu_rows, scale, out_f32, ld_out,
      (__nv_bfloat16*)out_split, out_parts, out_kp);
  FB_CHECK_LAUNCH("attention_relpos_simt");
  return FB200_OK;
}
}  // namespace fb200
