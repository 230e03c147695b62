// Dense contraction C = A . W^T over bf16 split operands on the 5th-gen tensor cores.
//
//   persistent CTAs (one per SM), warp-specialised:
//     warp 0      TMA producer   cp.async.bulk.tensor (128B swizzle) -> shared-memory stage ring
//     warp 1      MMA issuer     tcgen05.mma kind::f16, one elected thread, fp32 accumulators in TMEM
//     warps 2..5  epilogue       tcgen05.ld -> bias/ReLU/residual/bf16 split -> swizzled smem boxes -> TMA stores
//   two TMEM accumulator buffers (2 x BN columns) so the epilogue of tile i overlaps the MMAs of
//   tile i+1.
//
// Split operands: a stage holds ALL parts (hi, mid[, lo]) of the A and W k-block, and every cross
// term of the split is issued against that one stage -- the tiles are fetched once from L2 for
// 3 (or 6) tensor-core passes, which is what keeps the fp32-parity mode off the L2 roof.
//   NP = 1: bf16                      1 term
//   NP = 2: hi + mid (16-bit mant.)   3 terms  hi.mid, mid.hi, hi.hi        ("bf16x3")
//   NP = 3: hi + mid + lo             6 terms  + mid.mid, hi.lo, lo.hi
// A SIMT fp32 kernel over the SAME operands and epilogue is kept for bring-up tests (impl=1).
#include <cuda.h>

#include "common.cuh"
#include "sm100_ptx.cuh"
#include "stack.cuh"

namespace fb200 {

constexpr int BM = 128;
constexpr int BK = 64;  // bf16 elements = 128 bytes = one swizzle row
constexpr int kGemmThreads = 192;  // 2 + 4 epilogue warps; EW = 8 adds four more (64 + 32 * EW threads)
constexpr int kEpiWarpBytes = 8192;  // per epilogue warp: fp32 box 32x32 (4 KB, SW128) + 2 bf16 part boxes 32x32 (2 KB each, SW64)
// swish y * sigmoid(y) on the SFU (ex2.approx + rcp.approx: ~2 ulp, far inside the 16-bit-mantissa operand rounding that
// follows); the IEEE expf + division made the epilogue of the K = 256 FFN contractions longer than their mainloop
__device__ __forceinline__ float swish_fast(float y) { return __fdividef(y, 1.f + __expf(-y)); }

// y = acc (+bias)(relu)(+res1) for one output element (scalar path shared with the SIMT kernel)
__device__ __forceinline__ float epilogue_value(const GemmParams& p, int row, int col, float acc) {
  float y = acc;
  if (p.bias) y += __ldg(p.bias + col);
  if (p.act == 1) y = fmaxf(y, 0.f);
  if (p.act == 2) y = swish_fast(y);
  if (p.res1) y += p.res1[(size_t)row * p.ld_res1 + col];
  return y;
}
__device__ __noinline__ void epilogue_store_scalar(const GemmParams& p, int row, int col, float acc) {
  if (row >= p.m) return;
  if (col < p.n) {
    const float y = epilogue_value(p, row, col, acc);
    if (p.out_f32 && col >= p.f32_col0) p.out_f32[(size_t)row * p.ld_out + col - p.f32_col0] = y;
    if (p.out_split) store_split(p.out_split + (size_t)row * (p.out_parts * p.out_kp), p.out_kp, col, y, p.out_parts);
  } else if (p.out_split && col < p.out_kp) {  // zero padding columns of a split output
    store_split(p.out_split + (size_t)row * (p.out_parts * p.out_kp), p.out_kp, col, 0.f, p.out_parts);
  }
}

// Cross terms of the split, smallest products first (hi = 0, mid = 1, lo = 2).
template <int NP> __host__ __device__ constexpr int term_count() { return NP == 1 ? 1 : NP == 2 ? 3 : 6; }
template <int NP> __host__ __device__ constexpr int term_a(int t) {
  return NP == 1 ? 0 : NP == 2 ? (t == 1 ? 1 : 0) : (t == 0 ? 1 : t == 2 ? 2 : t == 4 ? 1 : 0);
}
template <int NP> __host__ __device__ constexpr int term_b(int t) {
  return NP == 1 ? 0 : NP == 2 ? (t == 0 ? 1 : 0) : (t == 0 ? 1 : t == 1 ? 2 : t == 3 ? 1 : 0);
}

#ifdef FB200_TRACE
__device__ long long g_gemm_trace[3][1024];
#define GTRACE(w, tag) do { if (blockIdx.x == 0 && gti < 511) { g_gemm_trace[w][2 * gti] = tag; g_gemm_trace[w][2 * gti + 1] = clock64(); ++gti; } } while (0)
#else
#define GTRACE(w, tag) do { } while (0)
#endif

// ---------------------------------------------------------------- tcgen05 kernel --------
// CL = 2: a CTA PAIR (two SMs of one TPC, tcgen05 cta_group::2) computes a [256 x BN] tile: each CTA holds its own
// 128 rows of A and HALF of the weight rows of a k-block, the leader (rank 0) issues one M = 256 MMA that reads
// both halves, and each CTA's tensor memory receives its 128 accumulator rows.  A CTA then takes in
// NP * (16 KB + BN / 2 * 128 B) per k-block instead of NP * (16 KB + BN * 128 B): the mainloop was bound by
// the ~41 B/clk an SM can take in, not by the tensor pipe -- and the smaller stage buys a third stage.
// EW = 8: EIGHT epilogue warps, two per tensor-memory lane group (warp % 4), each taking one half of the tile's columns.
// The epilogue of a K = 512 product is longer than its mainloop and LATENCY-bound per warp (one epilogue warp per
// scheduler: tcgen05.ld -> bias -> convert -> st.shared -> fence -> UTMASTG each wait for the previous step); two warps per
// scheduler overlap those chains.  Costs 32 KB more staging, i.e. one ring stage less.
// EPIB = 4096: the fp32 box and the split boxes of an epilogue warp share ONE 4 KB staging area -- for products with a single
// kind of output (or outputs in disjoint column ranges that never meet in a chunk); eight warps then fit the 32 KB of four.
template <int NP, int BN, int STAGES, int CL, int EW = 4, int EPIB = kEpiWarpBytes>
__global__ void __launch_bounds__(64 + 32 * EW, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                    const __grid_constant__ CUtensorMap map_f32, const __grid_constant__ CUtensorMap map_split,
                    const GemmParams p) {
  constexpr uint32_t kAPart = BM * BK * 2;
  constexpr uint32_t kBPart = (BN / CL) * BK * 2;  // this CTA's share of the weight rows
  constexpr uint32_t kStageBytes = NP * (kAPart + kBPart);
  constexpr uint32_t kEpiBytes = EW * EPIB;
  constexpr uint32_t kTmemCols = 2 * BN;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B needs 1024 B alignment
  const uint32_t epi_base = smem_base + STAGES * kStageBytes;
  const uint32_t bar_base = epi_base + kEpiBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 2 + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
  const uint32_t bias_sm = bar_base + 256u;  // the current tile's BN bias values (fp32), shared by the four epilogue warps

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // work unit = (group of CL row blocks, column block); cluster g takes units g, g + #clusters, ...; CTA rank r
  // of the pair owns row block group * CL + r (past the last row block it still loads / multiplies, nothing is
  // stored: TMA clips)
  const int cta_rank = CL > 1 ? (int)cluster_ctarank() : 0;
  const int unit0 = blockIdx.x / CL, unit_step = gridDim.x / CL;
  const int total_tiles = ((p.tiles_m + CL - 1) / CL) * p.tiles_n;
  const int k_iters = p.taps * p.k_blocks;
  constexpr uint16_t kClusterMask = (uint16_t)((1u << CL) - 1u);
  const bool leader = cta_rank == 0;
#ifdef FB200_TRACE
  int gti = 0;
#endif

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_f32) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_split) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < STAGES; ++s) {
        mbar_init(full_bar(s), 1);  // a pair uses the leader's: it counts the bytes of both CTAs' loads
        mbar_init(empty_bar(s), 1);
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(tfull_bar(a), 1);
        mbar_init(tempty_bar(a), EW * CL);  // one arrival per epilogue warp (of both CTAs, on the leader's)
      }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    if (CL == 1) {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(kTmemCols));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    } else {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(kTmemCols));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::);
    }
  }
  tcgen05_fence_before();
  if (CL > 1)
    cluster_sync_all();  // the peers' barriers are initialised before anything is multicast to them
  else
    __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
  // Everything above ran under the previous kernel's tail.  The trigger comes AFTER the tensor-memory allocation: a CTA
  // of the next kernel that lands on this SM must find this CTA's columns already taken (it would otherwise hold them while
  // blocked in its own pdl_wait, and a CTA of THIS grid still waiting for columns would never get them).
  pdl_trigger();
  pdl_wait();  // the previous kernels' outputs (A, the residual) are visible from here

  if (warp == 0) {
    if (elect_one()) {  // ===== TMA producer =====
      int it = 0;
      for (int tile = unit0; tile < total_tiles; tile += unit_step) {
        const int m0 = ((tile / p.tiles_n) * CL + cta_rank) * BM, n0 = (tile % p.tiles_n) * BN;
        int a_cg = 0, w_r0 = n0;
        if (p.group_n > 0) {
          const int g = n0 / p.group_n;
          a_cg = g * p.a_col_step;
          w_r0 = g * p.w_group_rows + (n0 - g * p.group_n);
        }
        for (int tap = 0; tap < p.taps; ++tap) {
          const int a_r = m0 + p.tap_row[tap], a_c = p.a_col0 + p.tap_col[tap] + a_cg;
          for (int kb = 0; kb < p.k_blocks; ++kb, ++it) {
            const int s = it % STAGES;
            const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
            mbar_wait(empty_bar(s), ph ^ 1u);
            GTRACE(0, 1);
            const uint32_t sa = smem_base + s * kStageBytes;
            if (CL == 1) {
              mbar_expect_tx(full_bar(s), kStageBytes);
#pragma unroll
              for (int pa = 0; pa < NP; ++pa)
                tma_load_2d(sa + pa * kAPart, &map_a, a_c + pa * p.kp_a + kb * BK, a_r, full_bar(s));
#pragma unroll
              for (int pb = 0; pb < NP; ++pb)
                tma_load_2d(sa + NP * kAPart + pb * kBPart, &map_w, pb * p.kp_w + kb * BK, tap * p.n + w_r0, full_bar(s));
            } else {  // own A rows and own half of the weight rows; the bytes are counted on the leader's barrier
              // Only the leader arms the barrier (for both CTAs' bytes): the peer's bytes may land first, the phase
              // still cannot complete before the leader's arrival, and the peer is spared a remote arrive per stage.
              const uint32_t lbar = leader ? full_bar(s) : mapa_shared(full_bar(s), 0);
              if (leader) mbar_expect_tx(full_bar(s), CL * kStageBytes);
#pragma unroll
              for (int pa = 0; pa < NP; ++pa)
                tma_load_2d_pair(sa + pa * kAPart, &map_a, a_c + pa * p.kp_a + kb * BK, a_r, lbar);
#pragma unroll
              for (int pb = 0; pb < NP; ++pb)
                tma_load_2d_pair(sa + NP * kAPart + pb * kBPart, &map_w, pb * p.kp_w + kb * BK,
                                 tap * p.n + w_r0 + cta_rank * (BN / CL), lbar);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (leader && elect_one()) {  // ===== MMA issuer (of the pair: the leader CTA) =====
      constexpr uint32_t idesc = umma_idesc_bf16(BM * CL, BN);
      int it = 0, tl = 0;
      for (int tile = unit0; tile < total_tiles; tile += unit_step, ++tl) {
        const int acc = tl & 1;
        mbar_wait(tempty_bar(acc), (((uint32_t)tl >> 1) & 1u) ^ 1u);  // epilogue drained this buffer
        GTRACE(1, 10);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BN);
        for (int j = 0; j < k_iters; ++j, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
          mbar_wait(full_bar(s), ph);
          GTRACE(1, 11);
          tcgen05_fence_after();
          const uint32_t sa = smem_base + s * kStageBytes;
          const uint32_t sb = sa + NP * kAPart;
#pragma unroll
          for (int t = 0; t < term_count<NP>(); ++t) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              const uint64_t da = umma_desc_sw128(sa + term_a<NP>(t) * kAPart + k * 32);
              const uint64_t db = umma_desc_sw128(sb + term_b<NP>(t) * kBPart + k * 32);
              if (CL == 1)
                umma_bf16(tmem_d, da, db, idesc, (j > 0 || t > 0 || k > 0) ? 1u : 0u);
              else
                umma_bf16_pair(tmem_d, da, db, idesc, (j > 0 || t > 0 || k > 0) ? 1u : 0u);
            }
          }
          // frees the smem stage once these MMAs retire -- in every CTA whose loads land in it
          if (CL == 1)
            umma_commit(empty_bar(s));
          else
            umma_commit_pair(empty_bar(s), kClusterMask);
          GTRACE(1, 12);
        }
        if (CL == 1)
          umma_commit(tfull_bar(acc));
        else
          umma_commit_pair(tfull_bar(acc), kClusterMask);  // each CTA's epilogue reads its own 128 rows
      }
    }
  } else {  // ===== epilogue warps 2..5: TMEM lane group = warp % 4, lane = output row =====
    // TMEM -> registers -> (bias, ReLU, residual, bf16 split) -> swizzled smem boxes -> TMA stores.
    // The TMA engine does the address generation, coalescing and edge clipping; the warp only
    // touches each value once, which keeps this issue-bound branch short.
    const int lg = warp & 3;
    // columns of the tile this warp handles: all of them (EW = 4) or one half (EW = 8: warps 2..5 the first, 6..9 the second)
    constexpr int kColsPerWarp = BN / (EW / 4);
    static_assert(EW == 4 || (EW == 8 && BN % 64 == 0), "eight epilogue warps split the tile's columns in two halves of whole chunks");
    const int c_lo = ((warp - 2) >> 2) * kColsPerWarp, c_hi = c_lo + kColsPerWarp;
    const uint32_t fbuf = epi_base + (uint32_t)(warp - 2) * EPIB;           // 32 rows x 128 B, SWIZZLE_128B
    const uint32_t sbuf = fbuf + (EPIB >= 8192 ? 4096u : 0u);               // per part: 32 rows x 64 B, SWIZZLE_64B
    const int m = p.m, n = p.n, f32_col0 = p.f32_col0, out_parts = p.out_parts, out_kp = p.out_kp;
    const float* __restrict__ bias = p.bias;
    const float* __restrict__ res1 = p.res1;
    const bool has_f32 = p.out_f32 != nullptr, has_split = p.out_split != nullptr;
    const float lo = p.act == 1 ? 0.f : -INFINITY;
    const bool swish = p.act == 2;
    const int ncols = has_split ? out_kp : n;
    const uint32_t sw128 = (uint32_t)(lane & 7), sw64 = (uint32_t)((lane >> 1) & 3);
    // Residual, coalesced: the warp reads the NEXT chunk's [32 rows x 32 columns] one row per instruction
    // (rv[i] = element (row i, column `lane`), a whole 128-byte line per load) while the current chunk is
    // processed, then turns it through the (otherwise unused) split staging box so that each lane gets its
    // own row.  Used when no split output is produced (the residual GEMMs write fp32 only).
    const bool direct = p.direct_store != 0;
    const bool res_coal = res1 != nullptr && !has_split && !direct;
    float rv[32];
    auto load_res = [&](int tile, int c) {
      const int r0 = ((tile / p.tiles_n) * CL + cta_rank) * BM + lg * 32, col = (tile % p.tiles_n) * BN + c + lane;
      const bool col_ok = col < n;
      const float* rp = res1 + (size_t)r0 * p.ld_res1 + col;
#pragma unroll
      for (int i = 0; i < 32; ++i) rv[i] = (col_ok && r0 + i < m) ? __ldg(rp + (size_t)i * p.ld_res1) : 0.f;
    };
    // Bias: the BN values of a tile's columns are staged ONCE per tile in shared memory by the 128 epilogue threads (fetched
    // into registers a whole tile early) and every chunk reads its 32 values with 8 broadcast LDS.128.  Fetched from global
    // memory where they are used -- 8 broadcast LDG.128 per chunk and warp -- they kept the load/store unit busy for ~300 of
    // the ~1350 cycles a chunk takes, and the epilogue, not the MMAs, bounds the K = 512 products (profiles/r02_gemm_epilogue.md).
    constexpr int kEpiThreads = EW * 32;
    constexpr int kBiasPerThread = BN / kEpiThreads > 0 ? BN / kEpiThreads : 1;
    const int et = (warp - 2) * 32 + lane;  // 0..kEpiThreads-1
    float bnext[kBiasPerThread];
    auto fetch_bias = [&](int tile) {
      const int n0 = (tile % p.tiles_n) * BN;
#pragma unroll
      for (int u = 0; u < kBiasPerThread; ++u) {
        const int cc = et * kBiasPerThread + u;
        bnext[u] = (bias != nullptr && cc < BN && n0 + cc < n) ? __ldg(bias + n0 + cc) : 0.f;
      }
    };
    auto epi_sync = [&]() { asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory"); };
    int tl = 0;
    if (unit0 < total_tiles) fetch_bias(unit0);
    if (res_coal && unit0 < total_tiles) load_res(unit0, c_lo);
    for (int tile = unit0; tile < total_tiles; tile += unit_step, ++tl) {
      const int acc = tl & 1;
      const int m0 = ((tile / p.tiles_n) * CL + cta_rank) * BM, n0 = (tile % p.tiles_n) * BN;
      if (bias != nullptr) {
        epi_sync();  // every warp has read the previous tile's values
        if (et * kBiasPerThread < BN) {
#pragma unroll
          for (int u = 0; u < kBiasPerThread; ++u)
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(bias_sm + (uint32_t)(et * kBiasPerThread + u) * 4u), "f"(bnext[u]) : "memory");
        }
        epi_sync();  // staged
        if (tile + unit_step < total_tiles) fetch_bias(tile + unit_step);  // the next tile's values: in flight under this tile
      }
      mbar_wait(tfull_bar(acc), ((uint32_t)tl >> 1) & 1u);
      if (warp == 2 && lane == 0) GTRACE(2, 20);
      tcgen05_fence_after();
      const int row = m0 + lg * 32 + lane;
      const bool row_ok = row < m;
      float am = -INFINITY, asum = 0.f;  // fused greedy pick: this row's running maximum / sum(exp(y - max)) / first column
      int ai = 0x7fffffff;
#pragma unroll 1
      for (int c = c_lo; c < c_hi; c += 32) {
        const int col0 = n0 + c;
        const bool live = col0 < ncols;    // warp-uniform
        const bool full = col0 + 32 <= n;  // warp-uniform: no column guard needed
        float y[32];
        {
          uint32_t r[32];
          tmem_ld32(tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(acc * BN + c), r);
#pragma unroll
          for (int j = 0; j < 32; ++j) y[j] = __uint_as_float(r[j]);
        }
        if (warp == 2 && lane == 0) GTRACE(2, 22);
        if (c + 32 >= c_hi) {  // this warp's columns of the accumulator are read: hand the TMEM buffer back to the MMA warp
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (CL == 1)
              mbar_arrive(tempty_bar(acc));
            else
              mbar_arrive_cluster(mapa_shared(tempty_bar(acc), 0));  // the leader issues the MMAs of both CTAs
          }
        }
        if (warp == 2 && lane == 0) GTRACE(2, 26);
        if (!live) continue;
        if (bias != nullptr) {  // columns >= n hold zeros in the staged tile
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float4 b4;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                         : "=f"(b4.x), "=f"(b4.y), "=f"(b4.z), "=f"(b4.w)
                         : "r"(bias_sm + (uint32_t)(c + 4 * j) * 4u)
                         : "memory");
            y[4 * j] += b4.x; y[4 * j + 1] += b4.y; y[4 * j + 2] += b4.z; y[4 * j + 3] += b4.w;
          }
        }
        if (warp == 2 && lane == 0) GTRACE(2, 27);
        if (p.amax_part != nullptr) {  // logits never leave the SM: streaming (max, first argmax, sum of exponentials)
          float cm = -INFINITY;
          int ci = 0;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if ((full || col0 + j < n) && y[j] > cm) {
              cm = y[j];
              ci = j;
            }
          if (cm > am) {
            asum *= __expf(am - cm);  // exp(-inf) = 0 on the first chunk
            am = cm;
            ai = col0 + ci;
          }
          float cs = 0.f;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (full || col0 + j < n) cs += __expf(y[j] - am);
          asum += cs;
          continue;
        }
#pragma unroll
        for (int j = 0; j < 32; ++j) y[j] = fmaxf(y[j], lo);
        if (swish) {
#pragma unroll
          for (int j = 0; j < 32; ++j) y[j] = swish_fast(y[j]);
        }
        if (res_coal) {
          if (EPIB < 8192) {  // shared staging area: the previous chunk's output box must have been read before it is reused
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncwarp();
          }
          // rv -> box (128-byte swizzle, like the fp32 output box) -> this lane's row
#pragma unroll
          for (int i = 0; i < 32; ++i)
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbuf + (uint32_t)i * 128u + ((((uint32_t)lane >> 2) ^ ((uint32_t)i & 7u)) * 16u) +
                                                         ((uint32_t)lane & 3u) * 4u),
                         "f"(rv[i])
                         : "memory");
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float4 q;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                         : "=f"(q.x), "=f"(q.y), "=f"(q.z), "=f"(q.w)
                         : "r"(sbuf + (uint32_t)lane * 128u + ((uint32_t)j ^ sw128) * 16u)
                         : "memory");
            y[4 * j] += q.x; y[4 * j + 1] += q.y; y[4 * j + 2] += q.z; y[4 * j + 3] += q.w;
          }
          __syncwarp();
          // next chunk of this tile that is live, else the first chunk of this CTA's next tile
          if (c + 32 < c_hi && col0 + 32 < ncols)
            load_res(tile, c + 32);
          else if (tile + unit_step < total_tiles)
            load_res(tile + unit_step, c_lo);
        } else if (res1 != nullptr && row_ok) {
          const float* rp = res1 + (size_t)row * p.ld_res1 + col0;
          if (full && (p.ld_res1 & 3) == 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 q = *(reinterpret_cast<const float4*>(rp) + j);
              y[4 * j] += q.x; y[4 * j + 1] += q.y; y[4 * j + 2] += q.z; y[4 * j + 3] += q.w;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (col0 + j < n) y[j] += rp[j];
          }
        }
        if (!full) {  // columns >= n: zero (they are the K padding of a split output; fp32 ones are clipped)
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (col0 + j >= n) y[j] = 0.f;
        }
        if (warp == 2 && lane == 0) GTRACE(2, 23);
        if (direct) {
          // Plain stores: the thread owns 32 consecutive columns of its row (one 128-byte line of fp32, 64 bytes per bf16
          // part).  No staging box, no cross-proxy fence, no wait for the previous boxes, no ~200-cycle UTMASTG per box.
          if (row_ok) {
            if (has_f32 && col0 >= f32_col0) {
              float* dst = p.out_f32 + (size_t)row * p.ld_out + (col0 - f32_col0);
              if (full) {
#pragma unroll
                for (int j = 0; j < 8; ++j) reinterpret_cast<float4*>(dst)[j] = make_float4(y[4 * j], y[4 * j + 1], y[4 * j + 2], y[4 * j + 3]);
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (col0 + j < n) dst[j] = y[j];
              }
            }
            if (has_split) {
              __nv_bfloat16* dst = p.out_split + (size_t)row * (out_parts * out_kp) + col0;
#pragma unroll
              for (int pp = 0; pp < 2; ++pp) {
                if (pp < out_parts) {
#pragma unroll
                  for (int q = 0; q < 4; ++q) {
                    uint32_t w[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                      const int j = q * 8 + u * 2;
                      const __nv_bfloat162 h = __floats2bfloat162_rn(y[j], y[j + 1]);
                      w[u] = *reinterpret_cast<const uint32_t*>(&h);
                      y[j] -= __low2float(h);
                      y[j + 1] -= __high2float(h);
                    }
                    reinterpret_cast<uint4*>(dst + (size_t)pp * out_kp)[q] = make_uint4(w[0], w[1], w[2], w[3]);
                  }
                }
              }
            }
          }
          continue;
        }
        // the boxes of the previous chunk must have been read by the TMA engine before they are overwritten
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
        if (warp == 2 && lane == 0) GTRACE(2, 24);
        const bool do_f32 = has_f32 && col0 >= f32_col0;
        if (do_f32) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(fbuf + (uint32_t)lane * 128u + ((uint32_t)j ^ sw128) * 16u),
                         "f"(y[4 * j]), "f"(y[4 * j + 1]), "f"(y[4 * j + 2]), "f"(y[4 * j + 3])
                         : "memory");
          if (EPIB < 8192 && has_split) {  // shared staging area and both outputs in this chunk: the fp32 box leaves first
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              tma_store_2d(&map_f32, fbuf, col0 - f32_col0, m0 + lg * 32);
              asm volatile("cp.async.bulk.commit_group;" ::: "memory");
              asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
            __syncwarp();
          }
        }
        if (has_split) {
#pragma unroll
          for (int pp = 0; pp < 2; ++pp) {
            if (pp < out_parts) {
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                uint32_t w[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                  const int j = q * 8 + u * 2;
                  const __nv_bfloat162 h = __floats2bfloat162_rn(y[j], y[j + 1]);
                  w[u] = *reinterpret_cast<const uint32_t*>(&h);
                  y[j] -= __low2float(h);  // residual feeds the next part
                  y[j + 1] -= __high2float(h);
                }
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sbuf + (uint32_t)pp * 2048u + (uint32_t)lane * 64u + ((uint32_t)q ^ sw64) * 16u),
                             "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3])
                             : "memory");
              }
            }
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (warp == 2 && lane == 0) GTRACE(2, 25);
        if (lane == 0) {
          const int r0 = m0 + lg * 32;
          if (do_f32 && !(EPIB < 8192 && has_split)) tma_store_2d(&map_f32, fbuf, col0 - f32_col0, r0);
          if (has_split) {
            tma_store_2d(&map_split, sbuf, col0, r0);
            if (out_parts > 1) tma_store_2d(&map_split, sbuf + 2048u, out_kp + col0, r0);
          }
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          if (warp == 2) GTRACE(2, 21);
        }
      }
      if (p.amax_part != nullptr && row_ok) {  // one (max, sum, first column) triple per row, column tile and column half
        constexpr int kHalves = EW / 4;
        p.amax_part[((size_t)row * p.tiles_n + (tile % p.tiles_n)) * kHalves + ((warp - 2) >> 2)] = make_float4(am, asum, __int_as_float(ai), 0.f);
      }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tcgen05_fence_before();
  if (CL > 1)
    cluster_sync_all();  // no CTA leaves while a peer may still multicast into it or arrive on its barriers
  else
    __syncthreads();
  if (warp == 1) {
    __syncwarp();
    if (CL == 1)
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
    else
      asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
  }
}

}  // namespace fb200

#ifdef FB200_TRACE
extern "C" __attribute__((visibility("default"))) int fb200_debug_gemm_trace(long long* out) {
  return (int)cudaMemcpyFromSymbol(out, fb200::g_gemm_trace, sizeof(long long) * 3072);
}
#endif

#include "gemm_host.inc"
