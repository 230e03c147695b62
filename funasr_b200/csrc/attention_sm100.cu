// Var-len multi-head attention on the 5th-gen tensor cores, streaming (online) softmax.
//
//   one CTA = 128 query rows of one (utterance, head); key/value tiles of 64 rows stream through
//   shared memory by TMA; the [T, T] score matrix of the reference
//   (funasr/models/sanm/attention.py:309) never exists.
//     warp 4      TMA producer + tcgen05.mma issuer (one elected thread)
//     warps 0..3  softmax: thread r owns query row r = TMEM lane r
//   per key tile j:   S_j = Q K_j^T          tcgen05, Q and K K-major (d contiguous)      -> TMEM
//                     P_j = exp2(S_j c - m)  registers, written as a bf16 split A operand  -> smem
//                     O_j = P_j V_j          tcgen05, V MN-major (as it lies in memory)    -> TMEM
//                     o   = o * corr + O_j   registers (fp32 running output, 128 per thread)
//   Operands are bf16 splits (hi [+ mid]) like the GEMM's: NP = 2 issues the 3 cross terms against
//   the same shared-memory tiles (fp32-parity mode), NP = 1 is plain bf16.
// A SIMT fp32 kernel over the SAME split operands is kept as impl = 1 (bring-up check, and the
// path for NP = 3 whose tiles do not fit in shared memory).
#include <math.h>

#include "common.cuh"
#include "sm100_ptx.cuh"

namespace fb200 {

constexpr int TQ = 128;   // query rows per CTA
constexpr int TK = 64;    // key/value rows per tile
constexpr int HD = 128;   // head dim
constexpr int kAttnThreads = 160;

struct AttnParams {
  int q_col0, q_ps, k_col0, v_col0, kv_ps;  // column offsets / part strides (elements)
  const int32_t* cu_q;
  const int32_t* cu_kv;
  float scale_log2e;  // d_k^-0.5 * log2(e)
  float* out_f32;
  int ld_out;
  __nv_bfloat16* out_split;
  int out_parts, out_kp;
  uint32_t v_lbo, v_sbo;
};

template <int NP>
__global__ void __launch_bounds__(kAttnThreads, 1)
attention_tcgen05_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                         const __grid_constant__ CUtensorMap map_v, const AttnParams p) {
  const int b = blockIdx.z, head = blockIdx.y;
  const int q_beg = p.cu_q[b], q_len = p.cu_q[b + 1] - q_beg;
  const int q0 = blockIdx.x * TQ;
  if (q0 >= q_len) return;
  const int kv_beg = p.cu_kv[b], kv_len = p.cu_kv[b + 1] - kv_beg;
  const int ntiles = (kv_len + TK - 1) / TK;

  constexpr uint32_t kQBlk = TQ * 128;  // one 64-column block of the Q tile (bytes)
  constexpr uint32_t kKBlk = TK * 128;
  constexpr uint32_t kQBytes = NP * 2 * kQBlk, kKBytes = NP * 2 * kKBlk, kPBytes = NP * TQ * 128;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t q_sm = base, k_sm = q_sm + kQBytes, v_sm = k_sm + kKBytes, p_sm = v_sm + kKBytes;
  const uint32_t bar_base = p_sm + kPBytes;
  const uint32_t bar_q = bar_base, bar_k = bar_base + 8, bar_v = bar_base + 16, bar_s = bar_base + 24,
                 bar_p = bar_base + 32, bar_o = bar_base + 40, tmem_slot = bar_base + 48;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 4) {
    if (lane == 0) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_q) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_k) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
      mbar_init(bar_q, 1);
      mbar_init(bar_k, 1);
      mbar_init(bar_v, 1);
      mbar_init(bar_s, 1);
      mbar_init(bar_p, 128);
      mbar_init(bar_o, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(256u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
  const uint32_t tmem_s = tmem_base, tmem_o = tmem_base + 128u;

  if (warp == 4) {
    if (lane == 0) {
      const int hcol = head * HD;
      auto load_k = [&](int j) {
        mbar_expect_tx(bar_k, kKBytes);
#pragma unroll
        for (int pp = 0; pp < NP; ++pp)
#pragma unroll
          for (int kb = 0; kb < 2; ++kb)
            tma_load_2d(k_sm + (pp * 2 + kb) * kKBlk, &map_k, p.k_col0 + pp * p.kv_ps + hcol + kb * 64,
                        kv_beg + j * TK, bar_k);
      };
      auto load_v = [&](int j) {
        mbar_expect_tx(bar_v, kKBytes);
#pragma unroll
        for (int pp = 0; pp < NP; ++pp)
#pragma unroll
          for (int db = 0; db < 2; ++db)
            tma_load_2d(v_sm + (pp * 2 + db) * kKBlk, &map_v, p.v_col0 + pp * p.kv_ps + hcol + db * 64,
                        kv_beg + j * TK, bar_v);
      };
      constexpr uint32_t idesc_s = umma_idesc_bf16(TQ, TK);
      constexpr uint32_t idesc_o = umma_idesc_bf16_bmn(TQ, HD);
      auto issue_s = [&]() {  // S = Q K^T, all cross terms of the split
        tcgen05_fence_after();
#pragma unroll
        for (int t = 0; t < (NP == 1 ? 1 : 3); ++t) {
          const int ta = NP == 1 ? 0 : (t == 1 ? 1 : 0), tb = NP == 1 ? 0 : (t == 0 ? 1 : 0);
#pragma unroll
          for (int kb = 0; kb < 2; ++kb)
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4)
              umma_bf16(tmem_s, umma_desc_sw128(q_sm + (ta * 2 + kb) * kQBlk + k4 * 32),
                        umma_desc_sw128(k_sm + (tb * 2 + kb) * kKBlk + k4 * 32), idesc_s, (t | kb | k4) ? 1u : 0u);
        }
        umma_commit(bar_s);
      };
      mbar_expect_tx(bar_q, kQBytes);
#pragma unroll
      for (int pp = 0; pp < NP; ++pp)
#pragma unroll
        for (int kb = 0; kb < 2; ++kb)
          tma_load_2d(q_sm + (pp * 2 + kb) * kQBlk, &map_q, p.q_col0 + pp * p.q_ps + hcol + kb * 64, q_beg + q0, bar_q);
      load_k(0);
      load_v(0);
      mbar_wait(bar_q, 0);
      mbar_wait(bar_k, 0);
      issue_s();
      for (int j = 0; j < ntiles; ++j) {
        const uint32_t ph = (uint32_t)j & 1u;
        mbar_wait(bar_s, ph);  // S_j done: the K tile is free
        if (j + 1 < ntiles) load_k(j + 1);
        mbar_wait(bar_p, ph);  // P_j in shared memory, S_j consumed
        mbar_wait(bar_v, ph);
        tcgen05_fence_after();
#pragma unroll
        for (int t = 0; t < (NP == 1 ? 1 : 3); ++t) {
          const int ta = NP == 1 ? 0 : (t == 1 ? 1 : 0), tb = NP == 1 ? 0 : (t == 0 ? 1 : 0);
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4)  // 16 key rows per MMA = two 8-row groups of the V tile
            umma_bf16(tmem_o, umma_desc_sw128(p_sm + ta * (TQ * 128) + k4 * 32),
                      umma_desc_sw128_mn(v_sm + tb * 2 * kKBlk + k4 * 2048, p.v_lbo, p.v_sbo), idesc_o,
                      (t | k4) ? 1u : 0u);
        }
        umma_commit(bar_o);
        if (j + 1 < ntiles) {
          mbar_wait(bar_k, ph ^ 1u);
          issue_s();  // overlaps the softmax warps' accumulation of O_j
        }
        mbar_wait(bar_o, ph);  // O_j done: the V and P tiles are free
        if (j + 1 < ntiles) load_v(j + 1);
      }
    }
  } else {
    const int r = warp * 32 + lane;  // query row within the tile = TMEM lane
    const uint32_t lane_off = (uint32_t)(warp * 32) << 16;
    float o[HD];
#pragma unroll
    for (int i = 0; i < HD; ++i) o[i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;
    const float c = p.scale_log2e;
    for (int j = 0; j < ntiles; ++j) {
      const uint32_t ph = (uint32_t)j & 1u;
      const int valid = min(TK, kv_len - j * TK);
      mbar_wait(bar_s, ph);
      tcgen05_fence_after();
      uint32_t s[32];
      float mx = -INFINITY;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        tmem_ld32(tmem_s + lane_off + h * 32, s);
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (h * 32 + i < valid) mx = fmaxf(mx, __uint_as_float(s[i]));
      }
      const float m_new = fmaxf(m_run, mx * c);
      const float corr = exp2f(m_run - m_new);  // exp2(-inf) = 0 on the first tile
      float sum = 0.f;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        tmem_ld32(tmem_s + lane_off + h * 32, s);
#pragma unroll
        for (int c8 = 0; c8 < 4; ++c8) {  // 8 key columns = one 16-byte chunk of the P row per part
          __nv_bfloat16 hi[8], mid[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int col = h * 32 + c8 * 8 + u;
            const float pv = col < valid ? exp2f(__uint_as_float(s[c8 * 8 + u]) * c - m_new) : 0.f;
            sum += pv;
            hi[u] = __float2bfloat16_rn(pv);
            mid[u] = __float2bfloat16_rn(pv - __bfloat162float(hi[u]));
          }
          const int chunk = h * 4 + c8;
          const uint32_t off = (uint32_t)(r >> 3) * 1024u + (uint32_t)(r & 7) * 128u + (uint32_t)((chunk ^ (r & 7)) * 16);
          const uint4 hv = *reinterpret_cast<uint4*>(hi);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(p_sm + off), "r"(hv.x), "r"(hv.y), "r"(hv.z), "r"(hv.w) : "memory");
          if (NP > 1) {
            const uint4 mv = *reinterpret_cast<uint4*>(mid);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(p_sm + (uint32_t)(TQ * 128) + off), "r"(mv.x), "r"(mv.y), "r"(mv.z), "r"(mv.w) : "memory");
          }
        }
      }
      l_run = l_run * corr + sum;
      m_run = m_new;
      fence_proxy_async_smem();  // make the P stores visible to the tensor core's (async proxy) reads
      tcgen05_fence_before();
      mbar_arrive(bar_p);
      mbar_wait(bar_o, ph);
      tcgen05_fence_after();
#pragma unroll
      for (int cb = 0; cb < HD; cb += 32) {
        tmem_ld32(tmem_o + lane_off + cb, s);
#pragma unroll
        for (int i = 0; i < 32; ++i) o[cb + i] = fmaf(o[cb + i], corr, __uint_as_float(s[i]));
      }
    }
    if (q0 + r < q_len) {
      const float inv = 1.f / l_run;
      const size_t row = (size_t)(q_beg + q0 + r);
      const int col0 = head * HD;
      if (p.out_f32) {
        float* dst = p.out_f32 + row * p.ld_out + col0;
#pragma unroll
        for (int i = 0; i < HD; i += 4)
          *reinterpret_cast<float4*>(dst + i) = make_float4(o[i] * inv, o[i + 1] * inv, o[i + 2] * inv, o[i + 3] * inv);
      }
      if (p.out_split) {
        __nv_bfloat16* dst = p.out_split + row * (size_t)(p.out_parts * p.out_kp);
#pragma unroll
        for (int i = 0; i < HD; i += 4) {
          const float y[4] = {o[i] * inv, o[i + 1] * inv, o[i + 2] * inv, o[i + 3] * inv};
          store_split4(dst, p.out_kp, col0 + i, y, p.out_parts);
        }
      }
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 4) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256u));
  }
}

}  // namespace fb200

#include "attention_host.inc"
