// Var-len multi-head attention with streaming (online) softmax, fp32 on the CUDA cores.
// One CTA = 64 query rows of one (utterance, head); K/V tiles of 64 rows stream through shared
// memory, so the [T, T] score matrix of the reference (funasr/models/sanm/attention.py:309) is
// never materialised.  This is the fp32-parity kernel; see DESIGN.md for the tcgen05 plan.
#include <math.h>

#include "common.cuh"

namespace fb200 {

constexpr int AQ = 64;    // query rows per CTA
constexpr int AK = 64;    // key/value rows per tile
constexpr int AD = 128;   // head dim
constexpr int ADP = AD + 4;
constexpr int APP = AK + 4;
constexpr int kAttnThreads = 256;
constexpr size_t kAttnSmem = (size_t)(AQ * ADP + AK * ADP + AK * AD + AQ * APP) * sizeof(float);

__global__ void __launch_bounds__(kAttnThreads)
attention_kernel(const float* __restrict__ q, int ld_q, const float* __restrict__ k, int ld_k,
                 const float* __restrict__ v, int ld_v, const int32_t* __restrict__ cu_q,
                 const int32_t* __restrict__ cu_kv, float scale, float* __restrict__ out_f32, int ld_out,
                 __nv_bfloat16* __restrict__ out_split, int nparts, int out_kp) {
  extern __shared__ float smem[];
  float* Qs = smem;                 // [AQ][ADP]
  float* Ks = Qs + AQ * ADP;        // [AK][ADP]
  float* Vs = Ks + AK * ADP;        // [AK][AD]
  float* Ps = Vs + AK * AD;         // [AQ][APP]

  const int b = blockIdx.z, head = blockIdx.y;
  const int q_beg = cu_q[b], q_len = cu_q[b + 1] - q_beg;
  const int q0 = blockIdx.x * AQ;
  if (q0 >= q_len) return;
  const int kv_beg = cu_kv[b], kv_len = cu_kv[b + 1] - kv_beg;
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int hoff = head * AD;

  // Q tile, pre-scaled like the reference (q_h * d_k ** -0.5 before the matmul)
  for (int i = tid; i < AQ * (AD / 4); i += kAttnThreads) {
    const int r = i / (AD / 4), c4 = i % (AD / 4);
    float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
    if (q0 + r < q_len) {
      val = *reinterpret_cast<const float4*>(q + (size_t)(q_beg + q0 + r) * ld_q + hoff + c4 * 4);
      val.x *= scale; val.y *= scale; val.z *= scale; val.w *= scale;
    }
    *reinterpret_cast<float4*>(Qs + r * ADP + c4 * 4) = val;
  }

  float o[4][8];
  float m_run[4], l_run[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    m_run[i] = -INFINITY;
    l_run[i] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) o[i][j] = 0.f;
  }

  for (int k0 = 0; k0 < kv_len; k0 += AK) {
    __syncthreads();  // previous tile fully consumed (also orders the Q stores on the first pass)
    for (int i = tid; i < AK * (AD / 4); i += kAttnThreads) {
      const int r = i / (AD / 4), c4 = i % (AD / 4);
      float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
      if (k0 + r < kv_len) {
        kv = *reinterpret_cast<const float4*>(k + (size_t)(kv_beg + k0 + r) * ld_k + hoff + c4 * 4);
        vv = *reinterpret_cast<const float4*>(v + (size_t)(kv_beg + k0 + r) * ld_v + hoff + c4 * 4);
      }
      *reinterpret_cast<float4*>(Ks + r * ADP + c4 * 4) = kv;
      *reinterpret_cast<float4*>(Vs + r * AD + c4 * 4) = vv;
    }
    __syncthreads();

    // S[ty*4+i][tx+16j] = Q . K
    float s[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) s[i][j] = 0.f;
#pragma unroll 4
    for (int d = 0; d < AD; d += 4) {
      float4 qv[4], kv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) qv[i] = *reinterpret_cast<const float4*>(Qs + (ty * 4 + i) * ADP + d);
#pragma unroll
      for (int j = 0; j < 4; ++j) kv[j] = *reinterpret_cast<const float4*>(Ks + (tx + 16 * j) * ADP + d);
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
    // online softmax; the 16 lanes sharing `ty` hold one row group
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (k0 + tx + 16 * j >= kv_len) s[i][j] = -INFINITY;
        mx = fmaxf(mx, s[i][j]);
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
      const float m_new = fmaxf(m_run[i], mx);
      const float corr = expf(m_run[i] - m_new);  // exp(-inf) = 0 on the first tile
      float sum = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float pv = expf(s[i][j] - m_new);
        sum += pv;
        Ps[(ty * 4 + i) * APP + tx + 16 * j] = pv;
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
      l_run[i] = l_run[i] * corr + sum;
      m_run[i] = m_new;
#pragma unroll
      for (int j = 0; j < 8; ++j) o[i][j] *= corr;
    }
    __syncthreads();
    // O[ty*4+i][tx*4 .. +3] and [64 + tx*4 .. +3] += P . V
#pragma unroll 2
    for (int kk = 0; kk < AK; kk += 4) {
      float4 pv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) pv[i] = *reinterpret_cast<const float4*>(Ps + (ty * 4 + i) * APP + kk);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float4 v0 = *reinterpret_cast<const float4*>(Vs + (kk + u) * AD + tx * 4);
        const float4 v1 = *reinterpret_cast<const float4*>(Vs + (kk + u) * AD + 64 + tx * 4);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float pu = u == 0 ? pv[i].x : u == 1 ? pv[i].y : u == 2 ? pv[i].z : pv[i].w;
          o[i][0] = fmaf(pu, v0.x, o[i][0]);
          o[i][1] = fmaf(pu, v0.y, o[i][1]);
          o[i][2] = fmaf(pu, v0.z, o[i][2]);
          o[i][3] = fmaf(pu, v0.w, o[i][3]);
          o[i][4] = fmaf(pu, v1.x, o[i][4]);
          o[i][5] = fmaf(pu, v1.y, o[i][5]);
          o[i][6] = fmaf(pu, v1.z, o[i][6]);
          o[i][7] = fmaf(pu, v1.w, o[i][7]);
        }
      }
    }
  }

#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = q0 + ty * 4 + i;
    if (r >= q_len) continue;
    const float inv = 1.f / l_run[i];
    const size_t row = (size_t)(q_beg + r);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int col = hoff + h * 64 + tx * 4;
      const float y0 = o[i][h * 4 + 0] * inv, y1 = o[i][h * 4 + 1] * inv, y2 = o[i][h * 4 + 2] * inv,
                  y3 = o[i][h * 4 + 3] * inv;
      if (out_f32) *reinterpret_cast<float4*>(out_f32 + row * ld_out + col) = make_float4(y0, y1, y2, y3);
      if (out_split) {
        __nv_bfloat16* srow = out_split + row * (size_t)(nparts * out_kp);
        store_split(srow, out_kp, col + 0, y0, nparts);
        store_split(srow, out_kp, col + 1, y1, nparts);
        store_split(srow, out_kp, col + 2, y2, nparts);
        store_split(srow, out_kp, col + 3, y3, nparts);
      }
    }
  }
}

}  // namespace fb200

using namespace fb200;

extern "C" __attribute__((visibility("default"))) int fb200_attention(const float* q, int ld_q, const float* k, int ld_k, const float* v, int ld_v,
                               const int32_t* cu_q, const int32_t* cu_kv, int batch, int q_rows, int max_q, int heads,
                               int head_dim, float scale, float* out_f32, int ld_out, void* out_split, int nparts,
                               void* stream) {
  FB_CHECK_ARG(head_dim == AD, "attention: head_dim %d unsupported (128 only)", head_dim);
  FB_CHECK_ARG(ld_q % 4 == 0 && ld_k % 4 == 0 && ld_v % 4 == 0, "attention: leading dims must be multiples of 4");
  FB_CHECK_ARG(out_f32 != nullptr || out_split != nullptr, "attention: no output");
  FB_CHECK_ARG(out_f32 == nullptr || ld_out % 4 == 0, "attention: ld_out must be a multiple of 4");
  FB_CHECK_ARG(out_split == nullptr || (nparts >= 1 && nparts <= 3), "attention: nparts must be 1, 2 or 3");
  FB_CHECK_ARG(batch > 0 && batch <= 65535 && heads > 0, "attention: batch %d / heads %d out of range", batch, heads);
  if (q_rows == 0 || max_q == 0) return FB200_OK;
  ProfScope prof(kProfAttention, stream, 0.0);
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kAttnSmem);
    if (e != cudaSuccess) {
      set_error("attention: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return FB200_ECUDA;
    }
    attr_set = true;
  }
  dim3 grid((max_q + AQ - 1) / AQ, heads, batch);
  attention_kernel<<<grid, kAttnThreads, kAttnSmem, (cudaStream_t)stream>>>(
      q, ld_q, k, ld_k, v, ld_v, cu_q, cu_kv, scale, out_f32, ld_out, (__nv_bfloat16*)out_split, nparts,
      round_up(heads * head_dim, 64));
  FB_CHECK_LAUNCH("attention");
  return FB200_OK;
}
