// Var-len multi-head attention on the 5th-gen tensor cores, streaming (online) softmax.
//
//   one CTA = 128 query rows of one (utterance, head); key/value tiles of 64 rows stream through
//   shared memory by TMA; the [T, T] score matrix of the reference
//   (funasr/models/sanm/attention.py:309) never exists.
//     warp 4      TMA producer + tcgen05.mma issuer (one elected thread)
//     warps 0..3  softmax: thread r owns query row r = TMEM lane r
//   per key tile j:   S_j = Q K_j^T          tcgen05, Q and K K-major (d contiguous)      -> TMEM (2 buffers)
//                     P_j = exp2(S_j c - m)  registers, written as a bf16 split A operand  -> smem (2 buffers)
//                     O  += P_j V_j          tcgen05, V MN-major (as it lies in memory)    -> TMEM (accumulates)
//   S, K and P are double-buffered so the tensor pipe runs S_{j+1} / P.V_{j-1} while the softmax
//   warps work on tile j.  O stays in TMEM for the whole row block; the running maximum `m` is only
//   raised (and O, l rescaled through tcgen05.ld/st) when a tile's maximum exceeds it by more than
//   2^8 -- softmax is invariant to the shift, so the result is unchanged (lazy rescale).
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
constexpr float kRescaleThreshold = 8.0f;  // log2 units

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
  const uint32_t q_sm = base, k_sm = q_sm + kQBytes, v_sm = k_sm + 2 * kKBytes, p_sm = v_sm + kKBytes;
  const uint32_t bar_base = p_sm + 2 * kPBytes;
  const uint32_t bar_q = bar_base, bar_v = bar_base + 8, bar_o = bar_base + 16, tmem_slot = bar_base + 24;
  auto bar_k = [&](int i) { return bar_base + 32u + 8u * i; };
  auto bar_s = [&](int i) { return bar_base + 48u + 8u * i; };
  auto bar_p = [&](int i) { return bar_base + 64u + 8u * i; };
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 4) {
    if (lane == 0) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_q) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_k) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
      mbar_init(bar_q, 1);
      mbar_init(bar_v, 1);
      mbar_init(bar_o, 1);
      for (int i = 0; i < 2; ++i) {
        mbar_init(bar_k(i), 1);
        mbar_init(bar_s(i), 1);
        mbar_init(bar_p(i), 128);
      }
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
  const uint32_t tmem_o = tmem_base + 128u;                  // O: columns [128, 256)
  auto tmem_s = [&](int i) { return tmem_base + 64u * i; };  // S buffers: columns [0, 64) and [64, 128)

  if (warp == 4) {
    if (lane == 0) {
      const int hcol = head * HD;
      auto load_k = [&](int j) {  // tile j -> K buffer j & 1
        const int i = j & 1;
        mbar_expect_tx(bar_k(i), kKBytes);
#pragma unroll
        for (int pp = 0; pp < NP; ++pp)
#pragma unroll
          for (int kb = 0; kb < 2; ++kb)
            tma_load_2d(k_sm + i * kKBytes + (pp * 2 + kb) * kKBlk, &map_k, p.k_col0 + pp * p.kv_ps + hcol + kb * 64,
                        kv_beg + j * TK, bar_k(i));
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
      auto issue_s = [&](int j) {  // S_j = Q K_j^T into S buffer j & 1, all cross terms of the split
        const int i = j & 1;
        mbar_wait(bar_k(i), ((uint32_t)j >> 1) & 1u);
        tcgen05_fence_after();
#pragma unroll
        for (int t = 0; t < (NP == 1 ? 1 : 3); ++t) {
          const int ta = NP == 1 ? 0 : (t == 1 ? 1 : 0), tb = NP == 1 ? 0 : (t == 0 ? 1 : 0);
#pragma unroll
          for (int kb = 0; kb < 2; ++kb)
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4)
              umma_bf16(tmem_s(i), umma_desc_sw128(q_sm + (ta * 2 + kb) * kQBlk + k4 * 32),
                        umma_desc_sw128(k_sm + i * kKBytes + (tb * 2 + kb) * kKBlk + k4 * 32), idesc_s,
                        (t | kb | k4) ? 1u : 0u);
        }
        umma_commit(bar_s(i));
      };
      mbar_expect_tx(bar_q, kQBytes);
#pragma unroll
      for (int pp = 0; pp < NP; ++pp)
#pragma unroll
        for (int kb = 0; kb < 2; ++kb)
          tma_load_2d(q_sm + (pp * 2 + kb) * kQBlk, &map_q, p.q_col0 + pp * p.q_ps + hcol + kb * 64, q_beg + q0, bar_q);
      load_k(0);
      if (ntiles > 1) load_k(1);
      load_v(0);
      mbar_wait(bar_q, 0);
      issue_s(0);
      if (ntiles > 1) issue_s(1);
      for (int j = 0; j < ntiles; ++j) {
        const int i = j & 1;
        const uint32_t ph2 = ((uint32_t)j >> 1) & 1u;
        mbar_wait(bar_s(i), ph2);  // S_j done: K buffer i is free for tile j + 2
        if (j + 2 < ntiles) load_k(j + 2);
        mbar_wait(bar_p(i), ph2);  // P_j in shared memory, S buffer i consumed, O rescaled if needed
        mbar_wait(bar_v, (uint32_t)j & 1u);
        tcgen05_fence_after();
#pragma unroll
        for (int t = 0; t < (NP == 1 ? 1 : 3); ++t) {
          const int ta = NP == 1 ? 0 : (t == 1 ? 1 : 0), tb = NP == 1 ? 0 : (t == 0 ? 1 : 0);
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4)  // 16 key rows per MMA = two 8-row groups of the V tile
            umma_bf16(tmem_o, umma_desc_sw128(p_sm + i * kPBytes + ta * (TQ * 128) + k4 * 32),
                      umma_desc_sw128_mn(v_sm + tb * 2 * kKBlk + k4 * 2048, p.v_lbo, p.v_sbo), idesc_o,
                      (j | t | k4) ? 1u : 0u);
        }
        umma_commit(bar_o);
        if (j + 2 < ntiles) issue_s(j + 2);  // queued behind P.V_j; overlaps the softmax of tile j + 1
        mbar_wait(bar_o, (uint32_t)j & 1u);  // P.V_j done: the V tile and P buffer i are free
        if (j + 1 < ntiles) load_v(j + 1);
      }
    }
  } else {
#include "attention_softmax.inc"
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
