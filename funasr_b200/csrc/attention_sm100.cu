// Var-len multi-head attention on the 5th-gen tensor cores, streaming (online) softmax.
//
// Work item = 128 query rows of one (utterance, head); a PERSISTENT grid (one CTA per SM) walks the
// item list of the plan (fb200_attention_plan) round-robin, and every role keeps running across item
// boundaries, so the loads / first scores of the next item overlap the tail of the current one.
// The [T, T] score matrix of the reference (funasr/models/sanm/attention.py:309) never exists.
//   warps 0..3  softmax        thread r owns query row r = TMEM lane r
//   warps 4..7  Q / output     next item's Q tile: TMA -> staging -> registers -> TMEM; O / l -> TMA stores
//   warps 8,11  TMA producers  K (3 stages) / V (2 stages) tiles of 64 rows, one elected thread each
//   warp  9     P.V issuer     tcgen05.mma, one elected thread
//   warp 10     S issuer       tcgen05.mma, one elected thread, two tiles ahead
// The register file is re-divided per warpgroup (setmaxnreg): 208 / 208 / 88 registers per thread (the CTA owns 3 x 168).
// Both A operands live in TENSOR MEMORY (row = lane, two bf16 per 32-bit column), so shared memory
// only feeds the B operands -- with A from shared memory the 128x64x16 S = Q K^T steps were bound by
// shared-memory bandwidth, not by the tensor pipe:
//     S_j = Q K_j^T          A = Q in TMEM, B = K tile (K-major)                    -> TMEM (2 buffers)
//     P_j = exp2(S_j c - m)  registers, packed to bf16 hi (+ mid) pairs            -> TMEM (2 buffers)
//     O  += P_j V_j          A = P in TMEM, B = V tile (MN-major, as it lies)      -> TMEM (accumulates)
// The tensor pipe runs S_{j+2} / P.V_j while the softmax warps work on tile j + 1.  O stays in TMEM
// for the whole item; the running maximum `m` is only raised (and O, l rescaled through
// tcgen05.ld/st) when a tile's maximum exceeds it by more than 2^8 -- softmax is invariant to the
// shift, so the result is unchanged (lazy rescale).
// Operands are bf16 splits (hi [+ mid]) like the GEMM's: NP = 2 issues the 3 cross terms against
// the same tiles (fp32-parity mode), NP = 1 is plain bf16.
// Tiles are numbered G = 0, 1, .. per CTA across items; every barrier completes once per tile (or
// once per item) and each waiter consumes its phases in order.
// A SIMT fp32 kernel over the SAME split operands is kept as impl = 1 (bring-up check, and the
// path for NP = 3).
#include <math.h>

#include "common.cuh"
#include "sm100_ptx.cuh"
#include "stack.cuh"

namespace fb200 {

constexpr int TQ = 128;   // query rows per item
constexpr int TK = 64;    // key/value rows per tile
constexpr int HD = 128;   // head dim
constexpr int VS = 2;     // shared-memory stages of the V tiles
// ... of the K tiles: 3 with single-part operands, 2 with hi + mid parts (32 KB per stage) -- the third stage's 32 KB hold the
// output boxes there, so that the Q staging area is free for the NEXT item's Q tile while an item's output is written
__host__ __device__ constexpr int attn_k_stages(int np) { return np == 1 ? 3 : 2; }
constexpr int kAttnThreads = 384;  // 3 warpgroups (register budgets are set per warpgroup)
constexpr float kRescaleThreshold = 8.0f;  // log2 units
// tensor-memory columns
constexpr uint32_t kTmemS = 0, kTmemP = 128, kTmemO = 256, kTmemQ = 384, kTmemCols = 512;
// staging areas of the Q / output warps: the Q tile of the next item (NP parts x 2 column blocks of [128 rows x 128 B]),
// then two [32 x 64] output boxes per warp (separate areas: the next item's Q is fetched a whole item ahead)
__host__ __device__ constexpr uint32_t attn_qstage_bytes(int np) { return np * 2 * TQ * 128; }
constexpr uint32_t kAttnOutBytes = 4 * 8192;
__host__ __device__ constexpr uint32_t attn_stage_bytes(int np) { return attn_qstage_bytes(np) + kAttnOutBytes; }

#ifdef FB200_TRACE
__device__ long long g_attn_trace[8][512];
#define TRACE(w, tag) do { if (trace_on && ti < 255) { g_attn_trace[w][2 * ti] = tag; g_attn_trace[w][2 * ti + 1] = clock64(); ++ti; } } while (0)
#else
#define TRACE(w, tag) do { } while (0)
#endif

struct AttnParams {
  const __nv_bfloat16* q;  // split matrix; head h of part p at column q_col0 + p * q_ps + h * HD
  int ld_q, q_col0, q_ps;
  int k_col0, v_col0, kv_ps;  // column offsets / part stride (elements) inside the K / V tensor maps
  const int32_t* plan;        // [0] = number of row blocks, then int4 {q_row, q_valid, kv_beg, kv_len} each
  int heads;
  float scale_log2e;  // d_k^-0.5 * log2(e)
  float* out_f32;
  int ld_out;
  __nv_bfloat16* out_split;
  int out_parts, out_kp;
  uint32_t v_lbo, v_sbo;
};

struct AttnItem {
  int q_row, q_valid, kv_beg, kv_len, ntiles, hcol;
};
__device__ __forceinline__ AttnItem load_item(const AttnParams& p, int k) {
  const int4 t = __ldg(reinterpret_cast<const int4*>(p.plan) + 1 + k / p.heads);
  AttnItem it;
  it.q_row = t.x;
  it.q_valid = t.y;
  it.kv_beg = t.z;
  it.kv_len = t.w;
  it.ntiles = (t.w + TK - 1) / TK;
  it.hcol = (k % p.heads) * HD;
  return it;
}
// Position of one role in this CTA's tile sequence: item ordinal n, tile j of it, running tile number G.
struct TileCursor {
  int k, n, j, G, n_items;
  bool valid;
  AttnItem it;
  __device__ __forceinline__ void init(const AttnParams& p, int items) {
    n_items = items;
    k = blockIdx.x;
    n = j = G = 0;
    valid = k < n_items;
    if (valid) it = load_item(p, k);
  }
  __device__ __forceinline__ bool last_tile() const { return j == it.ntiles - 1; }
  __device__ __forceinline__ void next(const AttnParams& p) {
    ++G;
    if (++j == it.ntiles) {
      j = 0;
      ++n;
      k += gridDim.x;
      valid = k < n_items;
      if (valid) it = load_item(p, k);
    }
  }
};

// Plan: every (utterance, 128-row query block) that has keys becomes one entry, in utterance order.
__global__ void attention_plan_kernel(const int32_t* __restrict__ cu_q, const int32_t* __restrict__ cu_kv, int batch,
                                      int32_t* __restrict__ plan, int capacity) {
  __shared__ int s_count[1024];
  const int tid = threadIdx.x, nth = blockDim.x;
  const int per = (batch + nth - 1) / nth;
  const int b0 = min(batch, tid * per), b1 = min(batch, b0 + per);
  int cnt = 0;
  for (int b = b0; b < b1; ++b) {
    const int ql = cu_q[b + 1] - cu_q[b], kl = cu_kv[b + 1] - cu_kv[b];
    if (ql > 0 && kl > 0) cnt += (ql + TQ - 1) / TQ;
  }
  s_count[tid] = cnt;
  __syncthreads();
  for (int off = 1; off < nth; off <<= 1) {  // inclusive scan
    const int v = tid >= off ? s_count[tid - off] : 0;
    __syncthreads();
    s_count[tid] += v;
    __syncthreads();
  }
  int pos = s_count[tid] - cnt;
  if (tid == nth - 1) {
    plan[0] = min(s_count[tid], capacity);
    plan[1] = s_count[tid];  // > capacity: the caller's buffer was too small -- the attention kernels trap on it
    plan[2] = plan[3] = 0;
  }
  int4* tiles = reinterpret_cast<int4*>(plan) + 1;
  for (int b = b0; b < b1; ++b) {
    const int qb = cu_q[b], ql = cu_q[b + 1] - qb, kb = cu_kv[b], kl = cu_kv[b + 1] - kb;
    if (ql <= 0 || kl <= 0) continue;
    for (int q0 = 0; q0 < ql; q0 += TQ, ++pos)
      if (pos < capacity) tiles[pos] = make_int4(qb + q0, min(TQ, ql - q0), kb, kl);
  }
}

template <int NP>
__global__ void __launch_bounds__(kAttnThreads, 1)
attention_tcgen05_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                         const __grid_constant__ CUtensorMap map_v, const __grid_constant__ CUtensorMap map_o,
                         const AttnParams p) {
  constexpr uint32_t kKBlk = TK * 128;           // one 64-column block of a K / V tile (bytes)
  constexpr uint32_t kKBytes = NP * 2 * kKBlk;   // one stage: all parts, both column blocks
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);
  constexpr int KS = attn_k_stages(NP);
  const uint32_t k_sm = base, v_sm = k_sm + KS * kKBytes, stage_sm = v_sm + VS * kKBytes;
  const uint32_t outbox_sm = stage_sm + attn_qstage_bytes(NP);
  const uint32_t l_sm = stage_sm + attn_stage_bytes(NP);  // row sums of four items
  const uint32_t bar_base = l_sm + 4 * TQ * 4;
  const uint32_t tmem_slot = bar_base;
  // Every barrier is built so that its next completion needs work only its waiters can release (full/empty
  // pairs), so no waiter can ever fall two phases behind and misread a parity.
  const uint32_t bar_q = bar_base + 8;    // Q of item n in tensor memory            (128 arrivals, per item)
  const uint32_t bar_qf = bar_base + 16;  // all S of item n done: Q may be replaced (commit, per item)
  const uint32_t bar_of = bar_base + 24;  // all P.V of item n done: O is final       (commit, per item)
  const uint32_t bar_oe = bar_base + 32;  // O of item n copied out of tensor memory  (128 arrivals, per item)
  auto bar_l = [&](int i) { return bar_base + 40u + 8u * i; };   // row sums of item n  (128 arrivals, n & 1)
  const uint32_t bar_qs = bar_base + 184;  // Q tile of the next item in the staging area (tx, per item)
  auto bar_k = [&](int i) { return bar_base + 56u + 8u * i; };   // K stage filled      (tx, per tile)
  auto bar_ke = [&](int i) { return bar_base + 80u + 8u * i; };  // K stage consumed    (commit, per tile)
  auto bar_v = [&](int i) { return bar_base + 104u + 8u * i; };  // V stage filled      (tx, per tile)
  auto bar_ve = [&](int i) { return bar_base + 120u + 8u * i; }; // V stage consumed    (commit, per tile)
  auto bar_s = [&](int i) { return bar_base + 136u + 8u * i; };  // S_G done            (commit, per tile, G & 1)
  auto bar_p = [&](int i) { return bar_base + 152u + 8u * i; };  // P_G written         (128 arrivals, per tile)
  auto bar_o = [&](int i) { return bar_base + 168u + 8u * i; };  // P.V_G done          (commit, per tile, G & 1)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#ifdef FB200_TRACE
  const bool trace_on = blockIdx.x == 0 && lane == 0 && (warp == 0 || warp == 4 || (warp >= 8 && warp <= 10));
  const int tw = warp == 0 ? 0 : warp == 4 ? 1 : warp == 8 ? 2 : warp == 10 ? 7 : 3;
  int ti = 0;
  TRACE(tw, 1);
#endif

  if (warp == 8) {
    if (elect_one()) {
      if (base & 1023u) {
        printf("fb200: attention shared memory base is not 1024-byte aligned\n");
        __trap();
      }
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_q) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_o) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_k) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
      mbar_init(bar_q, 128);
      mbar_init(bar_qf, 1);
      mbar_init(bar_of, 1);
      mbar_init(bar_oe, 128);
      mbar_init(bar_qs, 1);
      for (int i = 0; i < KS; ++i) {
        mbar_init(bar_k(i), 1);
        mbar_init(bar_ke(i), 1);
      }
      for (int i = 0; i < VS; ++i) {
        mbar_init(bar_v(i), 1);
        mbar_init(bar_ve(i), 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(bar_l(i), 128);
        mbar_init(bar_s(i), 1);
        mbar_init(bar_p(i), 128);
        mbar_init(bar_o(i), 1);
      }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
  } else if (warp == 9) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(kTmemCols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
  const uint32_t tmem_o = tmem_base + kTmemO, tmem_q = tmem_base + kTmemQ;
  auto tmem_s = [&](int i) { return tmem_base + kTmemS + 64u * i; };  // fp32 scores of a tile: 64 columns
  auto tmem_p = [&](int i) { return tmem_base + kTmemP + 64u * i; };  // bf16 pairs: 32 columns per part
  // Everything above (barriers, tensor memory, descriptor prefetch) ran under the previous kernel's tail; the plan, Q, K
  // and V are other kernels' outputs and are visible from here on.
  pdl_trigger();  // after the tensor-memory allocation (see gemm_sm100.cu): the next kernel's CTAs may be scheduled now
  pdl_wait();
  if (p.plan[1] > p.plan[0]) {  // the plan was cut at its capacity: query blocks would silently go missing -- fail loudly
    if (threadIdx.x == 0 && blockIdx.x == 0)
      printf("fb200_attention: the plan needs %d entries but was built with capacity %d (fb200_attention_plan: capacity >= "
             "q_rows / 128 + batch always suffices)\n", p.plan[1], p.plan[0]);
    __trap();
  }
  const int n_items = p.plan[0] * p.heads;
  const bool idle = (int)blockIdx.x >= n_items;  // CTA-uniform: nothing to do but hand the tensor memory back

  if (idle) {
  } else if (warp >= 8) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 88;");
    if (warp == 8 || warp == 11) {
    // ------------------------------------------------------------------ TMA producers -----
    // warp 8 streams the K tiles, warp 11 the V tiles (one elected thread each), so that neither waits for
    // the other's stages
    if (elect_one()) {
      TileCursor c;
      c.init(p, n_items);
      if (warp == 8) {
        while (c.valid) {
          const int G = c.G, i = G % KS;
          if (G >= KS) mbar_wait(bar_ke(i), (uint32_t)(G / KS - 1) & 1u);  // S_{G-KS} done: stage i is free
          mbar_expect_tx(bar_k(i), kKBytes);
#pragma unroll
          for (int pp = 0; pp < NP; ++pp)
#pragma unroll
            for (int kb = 0; kb < 2; ++kb)
              tma_load_2d(k_sm + i * kKBytes + (pp * 2 + kb) * kKBlk, &map_k, p.k_col0 + pp * p.kv_ps + c.it.hcol + kb * 64,
                          c.it.kv_beg + c.j * TK, bar_k(i));
          TRACE(2, 40);
          c.next(p);
        }
      } else {
        while (c.valid) {
          const int G = c.G, i = G % VS;
          if (G >= VS) mbar_wait(bar_ve(i), (uint32_t)(G / VS - 1) & 1u);  // P.V_{G-VS} done: stage i is free
          mbar_expect_tx(bar_v(i), kKBytes);
#pragma unroll
          for (int pp = 0; pp < NP; ++pp)
#pragma unroll
            for (int db = 0; db < 2; ++db)
              tma_load_2d(v_sm + i * kKBytes + (pp * 2 + db) * kKBlk, &map_v, p.v_col0 + pp * p.kv_ps + c.it.hcol + db * 64,
                          c.it.kv_beg + c.j * TK, bar_v(i));
          c.next(p);
        }
      }
    }
    } else if (warp == 9 || warp == 10) {
    // ------------------------------------------------------------------ MMA issuers -----
    // Two warps feed the one tensor pipe: warp 9 the P.V products, warp 10 the S = Q K^T products two tiles
    // ahead.  Both are released by the same event (P_G written), are independent of each other otherwise, and
    // the pipe's queue is shallow -- an issuing thread that stops to wait for a barrier lets it run dry, so the
    // barrier latencies of one warp are covered by the other warp's queued work.  The whole warp walks the tile
    // sequence; the waits of one step are spread over lanes 0..2 in one common polling loop (their latencies,
    // ~150 cycles each even when satisfied, overlap), then one elected lane issues.
    constexpr uint32_t idesc_s = umma_idesc_bf16(TQ, TK);
    constexpr uint32_t idesc_o = umma_idesc_bf16_bmn(TQ, HD);
#ifdef FB200_TRACE
    int tl = 0;
#endif
    auto wait_lanes = [&](uint32_t bar, uint32_t parity, bool enabled) {
      bool done = !enabled;
      uint32_t spins = 0;
      while (!__all_sync(0xffffffffu, done)) {
        if (!done) {
          done = mbar_try_wait(bar, parity) != 0;
#ifdef FB200_TRACE
          if (done && blockIdx.x == 0 && lane < 4 && tl < 255 && warp == 9) {
            g_attn_trace[4 + lane][2 * tl] = 60 + lane;
            g_attn_trace[4 + lane][2 * tl + 1] = clock64();
            ++tl;
          }
#endif
        }
        if (++spins == (1u << 26)) {
          if (!done) printf("fb200: attention MMA warp %d wait timed out (block %d lane %d)\n", warp, blockIdx.x, lane);
          __trap();
        }
      }
    };
    TileCursor c;
    c.init(p, n_items);
    if (warp == 10) {
      while (c.valid) {
        const int G = c.G, i = G % KS;
        {
          // lane 0: S buffer G & 1 consumed (P_{G-2} written)   lane 1: K stage filled   lane 2: Q of a new item
          uint32_t bar = bar_p(G & 1), parity = ((uint32_t)(G - 2) >> 1) & 1u;
          bool en = lane == 0 && G >= 2;
          if (lane == 1) bar = bar_k(i), parity = (uint32_t)(G / KS) & 1u, en = true;
          if (lane == 2) bar = bar_q, parity = (uint32_t)c.n & 1u, en = c.j == 0;
          wait_lanes(bar, parity, en);
        }
        tcgen05_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int t = 0; t < (NP == 1 ? 1 : 3); ++t) {
            const int ta = NP == 1 ? 0 : (t == 1 ? 1 : 0), tb = NP == 1 ? 0 : (t == 0 ? 1 : 0);
#pragma unroll
            for (int kb = 0; kb < 2; ++kb)
#pragma unroll
              for (int k4 = 0; k4 < 4; ++k4)  // 16 head-dim elements = 8 TMEM columns of Q
                umma_bf16_ts(tmem_s(G & 1), tmem_q + ta * 64 + kb * 32 + k4 * 8,
                             umma_desc_sw128(k_sm + i * kKBytes + (tb * 2 + kb) * kKBlk + k4 * 32), idesc_s,
                             (t | kb | k4) ? 1u : 0u);
          }
          umma_commit(bar_s(G & 1));
          umma_commit(bar_ke(i));
          if (c.last_tile()) umma_commit(bar_qf);
          TRACE(7, 14);
        }
        __syncwarp();
        c.next(p);
      }
    } else {
      while (c.valid) {
        const int G = c.G, i = G & 1, st = G % VS;
        {
          // lane 0: P_G in tensor memory (O rescaled)   lane 1: V stage filled
          // lane 2: previous item's O read out (first tile of an item)
          uint32_t bar = bar_p(i), parity = ((uint32_t)G >> 1) & 1u;
          bool en = lane == 0;
          if (lane == 1) bar = bar_v(st), parity = (uint32_t)(G / VS) & 1u, en = true;
          if (lane == 2) bar = bar_oe, parity = (uint32_t)(c.n - 1) & 1u, en = c.j == 0 && c.n > 0;
          wait_lanes(bar, parity, en);
        }
        TRACE(3, 11);
        tcgen05_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int t = 0; t < (NP == 1 ? 1 : 3); ++t) {
            const int ta = NP == 1 ? 0 : (t == 1 ? 1 : 0), tb = NP == 1 ? 0 : (t == 0 ? 1 : 0);
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4)  // 16 key rows = 8 TMEM columns of P = two 8-row groups of the V tile
              umma_bf16_ts(tmem_o, tmem_p(i) + ta * 32 + k4 * 8,
                           umma_desc_sw128_mn(v_sm + st * kKBytes + tb * 2 * kKBlk + k4 * 2048, p.v_lbo, p.v_sbo),
                           idesc_o, (c.j | t | k4) ? 1u : 0u);
          }
          umma_commit(bar_o(i));
          umma_commit(bar_ve(st));
          if (c.last_tile()) umma_commit(bar_of);
          TRACE(3, 13);
        }
        __syncwarp();
        c.next(p);
      }
    }
    }
  } else if (warp >= 4) {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 208;");
#include "attention_output.inc"
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 208;");
#include "attention_softmax.inc"
  }
  TRACE(tw, 99);
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 9) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
  }
}

}  // namespace fb200

#ifdef FB200_TRACE
extern "C" __attribute__((visibility("default"))) int fb200_debug_attn_trace(long long* out) {
  return (int)cudaMemcpyFromSymbol(out, fb200::g_attn_trace, sizeof(long long) * 4096);
}
#endif

#include "attention_host.inc"
