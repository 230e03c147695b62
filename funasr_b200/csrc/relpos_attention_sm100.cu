// Legacy relative-position attention of the Conformer encoder on the 5th-gen tensor cores (head dim 64).
//
//   scores[i, j] = ((q_i + u) . k_j + shift(bd)[i, j]) * scale,   bd[i, c] = (q_i + v) . p_c   (a tcgen05 GEMM upstream)
//   shift(bd)[i, j] = bd[i, T-1-i+j] (j <= i),  0 (j = i+1),  bd[i+1, j-i-2] (j >= i+2)
//   (LegacyRelPositionMultiHeadedAttention.rel_shift, funasr/models/transformer/attention.py:249-269, in closed form)
//
// One CTA = one item of the plan (fb200_attention_plan): 128 query rows of one (utterance, head); key / value tiles of
// 64 rows stream through a 3-stage shared-memory ring.
//   warps 0..3  softmax      thread r owns query row r = TMEM lane r: stages the shifted bias tile of its warp's 32 rows
//                            through shared memory one tile ahead (4-byte cp.async along the band: coalesced global
//                            reads, conflict-free row reads),
//                            online softmax, P packed to bf16 hi (+ mid) pairs INTO TENSOR MEMORY, O kept in registers
//   warp 4      TMA producer Q tile once, then K_j / V_j tiles (one elected thread)
//   warp 5      MMA issuer   S_j = Q K_j^T (A, B from shared memory) two tiles ahead; PV_j = P_j V_j (A = P in tensor
//                            memory, B = V tile MN-major as it lies); fp32 accumulators in TMEM, double-buffered
// The tensor pipe computes S_{j+1}, S_{j+2} and PV_{j-1} while the softmax warps work on tile j; PV_{j-1} is folded into
// the register accumulator (with the rescale factor of tile j-1) after P_j has been handed over.
// Operands are bf16 splits like the GEMM's: NP = 2 issues the 3 cross terms (fp32-parity mode), NP = 1 plain bf16.
#include <math.h>

#include "common.cuh"
#include "sm100_ptx.cuh"

namespace fb200 {

int relpos_attention_simt(const void* q, int ld_q, int q_col0, int q_part_stride, const void* k, int ld_k, int k_col0,
                          const void* v, int ld_v, int v_col0, int kv_part_stride, int nparts, const float* bd, int ld_bd,
                          long long bd_head_stride, const int32_t* cu_rows, int batch, int rows, int max_len, int heads,
                          int head_dim, float scale, float* out_f32, int ld_out, void* out_split, int out_parts,
                          void* stream);

namespace relpos {

constexpr int TQ = 128, TK = 64, HD = 64;
constexpr int NS = 3;                 // K/V stages
constexpr int kThreads = 192;
constexpr int kBiasPitch = TK + 1;    // floats per staged bias row: row reads by lane hit 32 different banks
constexpr uint32_t kTile = TK * 128;  // [64 rows x 64 bf16] = 8 KB, 128-byte swizzle
constexpr uint32_t kQTile = TQ * 128; // [128 rows x 64 bf16] = 16 KB
// tensor-memory columns: S (2 x 64 fp32), P (2 x 64: 32 per part), PV (2 x 64 fp32)
constexpr uint32_t kTmemS = 0, kTmemP = 128, kTmemPV = 256, kTmemCols = 512;

struct Params {
  const int32_t* plan;  // fb200_attention_plan layout: [0] = entries, then int4 {q_row, q_valid, kv_beg, kv_len}
  int heads;
  int q_col0, q_ps, k_col0, v_col0, kv_ps;
  const float* bd;
  int ld_bd;
  long long bd_head_stride;
  float scale_log2e;
  float* out_f32;
  int ld_out;
  __nv_bfloat16* out_split;
  int out_parts, out_kp;
};

__host__ __device__ constexpr size_t smem_bytes(int np) {
  return (size_t)np * kQTile + (size_t)NS * np * 2 * kTile + 2 * 4 * 32 * kBiasPitch * 4 + 256 + 1024;
}

template <int NP>
__global__ void __launch_bounds__(kThreads, 1)
relpos_attention_tc_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                           const __grid_constant__ CUtensorMap map_v, const Params p) {
  const int item = blockIdx.x / p.heads, head = blockIdx.x % p.heads;
  if (p.plan[1] > p.plan[0]) {  // truncated plan: fail loudly (see attention_sm100.cu)
    if (threadIdx.x == 0 && blockIdx.x == 0)
      printf("fb200_attention_relpos: the plan needs %d entries but was built with capacity %d\n", p.plan[1], p.plan[0]);
    __trap();
  }
  if (item >= p.plan[0]) return;
  const int4 it = __ldg(reinterpret_cast<const int4*>(p.plan) + 1 + item);
  const int q_row = it.x, q_valid = it.y, beg = it.z, T = it.w;
  const int ntiles = (T + TK - 1) / TK;
  const int hcol = head * HD;

  constexpr uint32_t kStage = NP * 2 * kTile;  // K parts then V parts
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t q_sm = base, kv_sm = q_sm + NP * kQTile, bias_sm = kv_sm + NS * kStage;
  const uint32_t bar_base = bias_sm + 2 * 4 * 32 * kBiasPitch * 4;
  const uint32_t tmem_slot = bar_base;
  const uint32_t bar_q = bar_base + 8;
  auto bar_full = [&](int i) { return bar_base + 16u + 8u * i; };
  auto bar_empty = [&](int i) { return bar_base + 40u + 8u * i; };
  auto bar_s = [&](int i) { return bar_base + 64u + 8u * i; };
  auto bar_p = [&](int i) { return bar_base + 80u + 8u * i; };
  auto bar_pv = [&](int i) { return bar_base + 96u + 8u * i; };
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 4) {
    if (elect_one()) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_q) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_k) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
      mbar_init(bar_q, 1);
      for (int i = 0; i < NS; ++i) {
        mbar_init(bar_full(i), 1);
        mbar_init(bar_empty(i), 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(bar_s(i), 1);
        mbar_init(bar_p(i), 128);
        mbar_init(bar_pv(i), 1);
      }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
  } else if (warp == 5) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(kTmemCols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
  auto tmem_s = [&](int i) { return tmem_base + kTmemS + 64u * i; };
  auto tmem_p = [&](int i) { return tmem_base + kTmemP + 64u * i; };
  auto tmem_pv = [&](int i) { return tmem_base + kTmemPV + 64u * i; };

  if (warp == 4) {
    // ------------------------------------------------------------------ TMA producer -----
    if (elect_one()) {
      mbar_expect_tx(bar_q, NP * kQTile);
#pragma unroll
      for (int pp = 0; pp < NP; ++pp) tma_load_2d(q_sm + pp * kQTile, &map_q, p.q_col0 + pp * p.q_ps + hcol, q_row, bar_q);
      for (int j = 0; j < ntiles; ++j) {
        const int i = j % NS;
        if (j >= NS) mbar_wait(bar_empty(i), (uint32_t)(j / NS - 1) & 1u);  // PV_{j-NS} done: the stage is free
        mbar_expect_tx(bar_full(i), kStage);
#pragma unroll
        for (int pp = 0; pp < NP; ++pp) {
          tma_load_2d(kv_sm + i * kStage + pp * kTile, &map_k, p.k_col0 + pp * p.kv_ps + hcol, beg + j * TK, bar_full(i));
          tma_load_2d(kv_sm + i * kStage + (NP + pp) * kTile, &map_v, p.v_col0 + pp * p.kv_ps + hcol, beg + j * TK, bar_full(i));
        }
      }
    }
  } else if (warp == 5) {
    // ------------------------------------------------------------------ MMA issuer -----
    if (elect_one()) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(TQ, TK);
      constexpr uint32_t idesc_o = umma_idesc_bf16_bmn(TQ, HD);
      auto issue_s = [&](int j) {  // S_j = Q K_j^T into buffer j & 1
        const int i = j % NS;
        mbar_wait(bar_full(i), (uint32_t)(j / NS) & 1u);
        tcgen05_fence_after();
#pragma unroll
        for (int t = 0; t < (NP == 1 ? 1 : 3); ++t) {
          const int ta = NP == 1 ? 0 : (t == 1 ? 1 : 0), tb = NP == 1 ? 0 : (t == 0 ? 1 : 0);
#pragma unroll
          for (int k4 = 0; k4 < HD / 16; ++k4)
            umma_bf16(tmem_s(j & 1), umma_desc_sw128(q_sm + ta * kQTile + k4 * 32),
                      umma_desc_sw128(kv_sm + i * kStage + tb * kTile + k4 * 32), idesc_s, (t | k4) ? 1u : 0u);
        }
        umma_commit(bar_s(j & 1));
      };
      mbar_wait(bar_q, 0);
      issue_s(0);
      if (ntiles > 1) issue_s(1);
      for (int j = 0; j < ntiles; ++j) {
        const int i = j % NS;
        mbar_wait(bar_p(j & 1), ((uint32_t)j >> 1) & 1u);  // P_j in tensor memory (and S_j, PV_{j-2} consumed)
        tcgen05_fence_after();
#pragma unroll
        for (int t = 0; t < (NP == 1 ? 1 : 3); ++t) {
          const int ta = NP == 1 ? 0 : (t == 1 ? 1 : 0), tb = NP == 1 ? 0 : (t == 0 ? 1 : 0);
#pragma unroll
          for (int k4 = 0; k4 < TK / 16; ++k4)  // 16 key rows = 8 TMEM columns of P = two 8-row groups of the V tile
            umma_bf16_ts(tmem_pv(j & 1), tmem_p(j & 1) + ta * 32 + k4 * 8,
                         umma_desc_sw128_mn(kv_sm + i * kStage + (NP + tb) * kTile + k4 * 2048, kTile, 1024), idesc_o,
                         (t | k4) ? 1u : 0u);
        }
        umma_commit(bar_pv(j & 1));
        umma_commit(bar_empty(i));
        if (j + 2 < ntiles) issue_s(j + 2);
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax warps -----
    const int r = warp * 32 + lane;
    const uint32_t lane_off = (uint32_t)(warp * 32) << 16;
    const float c = p.scale_log2e;
    const float* bdh = p.bd + (size_t)head * p.bd_head_stride;
    const int qi0 = q_row - beg + warp * 32;  // utterance-local index of this warp's first query row
    const float* bias_all = reinterpret_cast<const float*>(smem_raw + (bias_sm - smem_u32(smem_raw)));
    float o[HD];
#pragma unroll
    for (int u = 0; u < HD; ++u) o[u] = 0.f;
    float m_run = -INFINITY, l_run = 0.f, corr_prev = 1.f;

    auto fold_pv = [&](int jj, float corr) {  // o = o * corr + PV_jj
      mbar_wait(bar_pv(jj & 1), ((uint32_t)jj >> 1) & 1u);
      tcgen05_fence_after();
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t t[32];
        tmem_ld32(tmem_pv(jj & 1) + lane_off + h * 32, t);
#pragma unroll
        for (int u = 0; u < 32; ++u) o[h * 32 + u] = fmaf(o[h * 32 + u], corr, __uint_as_float(t[u]));
      }
    };

    // Shifted bias tile of this warp's 32 rows, tile j -> buffer j & 1: lanes run along the band (coalesced), one row
    // at a time, as 4-byte cp.async copies (zero-filled outside the band) -- issued one tile ahead, no registers held.
    auto stage_bias = [&](int j) {
      const int kj0 = j * TK;
      const uint32_t dst0 = bias_sm + (uint32_t)(((j & 1) * 4 + warp) * 32 * kBiasPitch + lane) * 4u;
#pragma unroll 8
      for (int rr = 0; rr < 32; ++rr) {
        const int qi = qi0 + rr;
        const float* row0 = bdh + (size_t)(beg + qi) * p.ld_bd + (T - 1 - qi);  // [kj] for kj <= qi
        const float* row1 = bdh + (size_t)(beg + qi + 1) * p.ld_bd - (qi + 2);   // [kj] for kj >= qi + 2
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int kj = kj0 + h * 32 + lane;
          const bool in = qi < T && kj < T && kj != qi + 1;
          const float* src = in ? (kj <= qi ? row0 + kj : row1 + kj) : bdh;
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst0 + (uint32_t)(rr * kBiasPitch + h * 32) * 4u),
                       "l"(src), "r"(in ? 4 : 0)
                       : "memory");
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    stage_bias(0);

    for (int j = 0; j < ntiles; ++j) {
      const int kj0 = j * TK;
      const float* bias_w = bias_all + ((j & 1) * 4 + warp) * 32 * kBiasPitch;
      __syncwarp();  // every lane has read buffer (j + 1) & 1 (tile j - 1)
      if (j + 1 < ntiles) {
        stage_bias(j + 1);
        asm volatile("cp.async.wait_group 1;" ::: "memory");
      } else {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
      }
      __syncwarp();
      uint32_t s[TK], pm[TK / 2];
      mbar_wait(bar_s(j & 1), ((uint32_t)j >> 1) & 1u);
      tcgen05_fence_after();
      tmem_ld32_nowait(tmem_s(j & 1) + lane_off, s);
      tmem_ld32_nowait(tmem_s(j & 1) + lane_off + 32, s + 32);
      tmem_ld_wait();
      tmem_ld_fence32(s);
      tmem_ld_fence32(s + 32);
      const int valid = min(TK, T - kj0);
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int u = 0; u < TK; ++u) {
        float x = __uint_as_float(s[u]) + bias_w[lane * kBiasPitch + u];
        if (u >= valid) x = -INFINITY;
        s[u] = __float_as_uint(x);
        mx4[u & 3] = fmaxf(mx4[u & 3], x);
      }
      const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * c;
      const float m_new = fmaxf(m_run, mx);
      const float corr = exp2f(m_run - m_new);  // first tile: exp2(-inf) = 0
      const float neg_m = -m_new;
      float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
      for (int u = 0; u < TK / 2; ++u) {
        const float p0 = ex2_approx(fmaf(__uint_as_float(s[2 * u]), c, neg_m));
        const float p1 = ex2_approx(fmaf(__uint_as_float(s[2 * u + 1]), c, neg_m));
        sum0 += p0;
        sum1 += p1;
        const __nv_bfloat162 hh = __floats2bfloat162_rn(p0, p1);
        const uint32_t hi = *reinterpret_cast<const uint32_t*>(&hh);
        s[u] = hi;
        if (NP > 1) {
          const __nv_bfloat162 mm = __floats2bfloat162_rn(p0 - __uint_as_float(hi << 16), p1 - __uint_as_float(hi & 0xffff0000u));
          pm[u] = *reinterpret_cast<const uint32_t*>(&mm);
        }
      }
      l_run = l_run * corr + (sum0 + sum1);
      m_run = m_new;
      // P buffer j & 1 was last read by PV_{j-2}, whose completion this thread waited for in iteration j - 1
      tmem_st32_nowait(tmem_p(j & 1) + lane_off, s);
      if (NP > 1) tmem_st32_nowait(tmem_p(j & 1) + lane_off + 32, pm);
      tmem_st_wait();
      tcgen05_fence_before();
      mbar_arrive(bar_p(j & 1));
      if (j >= 1) fold_pv(j - 1, corr_prev);
      corr_prev = corr;
    }
    fold_pv(ntiles - 1, corr_prev);

    if (r < q_valid) {
      const float inv = 1.f / l_run;
      const size_t row = (size_t)(q_row + r);
      if (p.out_f32) {
        float4* op = reinterpret_cast<float4*>(p.out_f32 + row * p.ld_out + hcol);
#pragma unroll
        for (int u = 0; u < HD; u += 4) op[u >> 2] = make_float4(o[u] * inv, o[u + 1] * inv, o[u + 2] * inv, o[u + 3] * inv);
      }
      if (p.out_split) {
        __nv_bfloat16* srow = p.out_split + row * (size_t)(p.out_parts * p.out_kp) + hcol;
#pragma unroll
        for (int u = 0; u < HD; u += 8) {
          uint32_t hi[4], mid[4], lo[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float y0 = o[u + 2 * e] * inv, y1 = o[u + 2 * e + 1] * inv;
            const __nv_bfloat162 h = __floats2bfloat162_rn(y0, y1);
            hi[e] = *reinterpret_cast<const uint32_t*>(&h);
            const float r0 = y0 - __low2float(h), r1 = y1 - __high2float(h);
            const __nv_bfloat162 m = __floats2bfloat162_rn(r0, r1);
            mid[e] = *reinterpret_cast<const uint32_t*>(&m);
            const __nv_bfloat162 l = __floats2bfloat162_rn(r0 - __low2float(m), r1 - __high2float(m));
            lo[e] = *reinterpret_cast<const uint32_t*>(&l);
          }
          *reinterpret_cast<uint4*>(srow + u) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
          if (p.out_parts > 1) *reinterpret_cast<uint4*>(srow + p.out_kp + u) = make_uint4(mid[0], mid[1], mid[2], mid[3]);
          if (p.out_parts > 2) *reinterpret_cast<uint4*>(srow + 2 * p.out_kp + u) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
      }
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 5) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
  }
}

template <int NP>
static int launch(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const Params& p, int grid, cudaStream_t st) {
  constexpr size_t smem = smem_bytes(NP);
  static_assert(smem <= 232448, "shared memory budget");
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(relpos_attention_tc_kernel<NP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
      set_error("attention_relpos: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return FB200_ECUDA;
    }
    attr_set = true;
  }
  relpos_attention_tc_kernel<NP><<<grid, kThreads, smem, st>>>(mq, mk, mv, p);
  return FB200_OK;
}

}  // namespace relpos
}  // namespace fb200

using namespace fb200;

extern "C" __attribute__((visibility("default"))) int fb200_attention_relpos(
    const void* q, int ld_q, int q_col0, int q_part_stride, const void* k, int ld_k, int k_col0, const void* v, int ld_v,
    int v_col0, int kv_part_stride, int nparts, const float* bd, int ld_bd, long long bd_head_stride, const int32_t* cu_rows,
    const int32_t* plan, int batch, int rows, int max_len, int heads, int head_dim, float scale, float* out_f32, int ld_out,
    void* out_split, int out_parts, int impl, void* stream) {
  FB_CHECK_ARG(impl == 0 || impl == 1, "attention_relpos: impl %d unknown", impl);
  if (impl == 1 || nparts == 3)
    return relpos_attention_simt(q, ld_q, q_col0, q_part_stride, k, ld_k, k_col0, v, ld_v, v_col0, kv_part_stride, nparts, bd,
                                 ld_bd, bd_head_stride, cu_rows, batch, rows, max_len, heads, head_dim, scale, out_f32, ld_out,
                                 out_split, out_parts, stream);
  FB_CHECK_ARG(head_dim == relpos::HD, "attention_relpos: head_dim %d unsupported (64 only)", head_dim);
  FB_CHECK_ARG(nparts == 1 || nparts == 2, "attention_relpos: nparts must be 1 or 2 on the tensor-core path");
  FB_CHECK_ARG(plan != nullptr, "attention_relpos: the tensor-core path needs a plan (fb200_attention_plan over cu_rows, cu_rows)");
  FB_CHECK_ARG(ld_q % 8 == 0 && ld_k % 8 == 0 && ld_v % 8 == 0 && q_col0 % 8 == 0 && k_col0 % 8 == 0 && v_col0 % 8 == 0 &&
                   q_part_stride % 8 == 0 && kv_part_stride % 8 == 0,
               "attention_relpos: leading dims / offsets must be multiples of 8");
  FB_CHECK_ARG(out_f32 != nullptr || out_split != nullptr, "attention_relpos: no output");
  FB_CHECK_ARG(out_f32 == nullptr || ld_out % 4 == 0, "attention_relpos: ld_out must be a multiple of 4");
  FB_CHECK_ARG(out_split == nullptr || (out_parts >= 1 && out_parts <= 3), "attention_relpos: out_parts must be 1, 2 or 3");
  FB_CHECK_ARG(batch > 0 && heads > 0 && ld_bd >= max_len, "attention_relpos: batch %d / heads %d / ld_bd %d < %d", batch, heads,
               ld_bd, max_len);
  if (rows == 0 || max_len == 0) return FB200_OK;
  ProfScope prof(kProfRelAttention, stream, 4.0 * (double)rows * max_len * head_dim * heads);
  CUtensorMap mq, mk, mv;
  int rc = make_tma_map_bf16(&mq, q, (uint64_t)rows, (uint64_t)ld_q, relpos::TQ);
  if (rc) return rc;
  rc = make_tma_map_bf16(&mk, k, (uint64_t)rows, (uint64_t)ld_k, relpos::TK);
  if (rc) return rc;
  rc = make_tma_map_bf16(&mv, v, (uint64_t)rows, (uint64_t)ld_v, relpos::TK);
  if (rc) return rc;
  relpos::Params p;
  p.plan = plan;
  p.heads = heads;
  p.q_col0 = q_col0;
  p.q_ps = q_part_stride;
  p.k_col0 = k_col0;
  p.v_col0 = v_col0;
  p.kv_ps = kv_part_stride;
  p.bd = bd;
  p.ld_bd = ld_bd;
  p.bd_head_stride = bd_head_stride;
  p.scale_log2e = scale * 1.4426950408889634f;
  p.out_f32 = out_f32;
  p.ld_out = ld_out;
  p.out_split = (__nv_bfloat16*)out_split;
  p.out_parts = out_split ? out_parts : 0;
  p.out_kp = round_up(heads * head_dim, 64);
  const long long max_items = ((long long)rows / relpos::TQ + batch) * heads;  // the plan's capacity bound
  rc = nparts == 1 ? relpos::launch<1>(mq, mk, mv, p, (int)max_items, (cudaStream_t)stream)
                   : relpos::launch<2>(mq, mk, mv, p, (int)max_items, (cudaStream_t)stream);
  if (rc) return rc;
  FB_CHECK_LAUNCH("attention_relpos_tcgen05");
  return FB200_OK;
}
