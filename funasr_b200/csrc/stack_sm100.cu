// Request-stack kernel: a whole stretch of the path (the 50 SAN-M encoder layers, the 16 decoder layers) for a SMALL
// request -- one or a few utterances, at most a few hundred packed rows -- as ONE persistent cooperative kernel.
//
// Why: at 167 rows (one 10 s utterance, BASELINE configs[0]) every kernel of the per-op path sits at its fixed cost
// (launch, tensor-memory allocation, barrier init, descriptor fetch, pipeline fill and drain: ~10 us on the device per
// launch, 578 launches per request) while the work itself is ~0.1 ms of tensor time and 0.13 ms of weight streaming.
// Here the fixed costs are paid once: the kernel is resident on every SM (one CTA per SM, cooperative launch), keeps its
// tensor memory, its barriers and its shared-memory rings for the whole program, and walks an OP LIST
//     GEMM | LayerNorm | FSMN | attention
// recorded from the very same host calls the per-op path makes (fb200_gemm / fb200_layernorm / fb200_fsmn /
// fb200_attention between fb200_stack_begin and fb200_stack_end).  Dependent ops are separated by a device-wide barrier
// (one release/acquire counter in global memory); independent neighbours (FSMN and attention of a layer) are not -- the
// recorder derives that from the byte ranges each op reads and writes.
//
// Arithmetic: the GEMM, LayerNorm and FSMN ops compute bit for bit what the per-op kernels compute (same bf16 split
// operands, same term order per k-block, same epilogue, same two-pass LayerNorm, same tap order) -- the tile shapes
// differ, the sums do not.  Attention is a streaming-softmax fp32 SIMT kernel over (utterance, head, <= 8 query rows)
// items -- at this size spreading a few query rows over every SM beats eight 128-row tensor-core items -- so its
// results differ from the tcgen05 attention kernel's in the last bits (both are checked against the oracle).
//
// Small-M tiling: an SM takes in ~41 B/clk, so what bounds a phase is the bytes ONE CTA must ingest: the A tile
// [bm x K] hi+mid and the W tile [bn x K] hi+mid.  bm (32..128 rows, in 32-row TMA boxes) and bn (32 / 64 columns) are
// picked on the device per op from the ACTUAL row count so that the tiles cover the SMs in one round with the smallest
// per-CTA ingest; programs are recorded once for a row CAPACITY and run with the request's row counts (`dims`).
// The next GEMM's weights are prefetched into L2 (cp.async.bulk.prefetch.L2) while the current op runs.
#include <cuda.h>
#include <math.h>
#include <stdlib.h>

#include <vector>

#include "common.cuh"
#include "sm100_ptx.cuh"
#include "stack.cuh"

namespace fb200 {

constexpr int kStackThreads = 256;
constexpr int SBK = 64;
constexpr int kStStages = 4;
constexpr uint32_t kStAPart = 128 * 128;     // 128 rows x 128 B
constexpr uint32_t kStBPart = 64 * 128;      // at most 64 weight rows
constexpr uint32_t kStStageBytes = 2 * (kStAPart + kStBPart);
constexpr uint32_t kStRingBytes = kStStages * kStStageBytes;  // 196608
constexpr uint32_t kStEpiWarpBytes = 8192;
constexpr uint32_t kStEpiBytes = 4 * kStEpiWarpBytes;
constexpr uint32_t kStMiscBytes = 1024;      // barriers, tensor-memory slot, bias tile
constexpr uint32_t kStOpBytes = 256;
constexpr uint32_t kStSmemBytes = kStRingBytes + kStEpiBytes + kStMiscBytes + kStOpBytes + 1024;
constexpr int kStTmemCols = 128;             // two accumulator buffers of 64 columns
constexpr int kStackMaxDims = 8;

// attention overlay on the ring + epilogue staging (229376 bytes)
constexpr int kAtTile = 64;                         // keys per tile
constexpr uint32_t kAtKRow = 272;                   // 128 bf16 + 16 B: conflict-free LDS.128 across keys
constexpr uint32_t kAtKPart = kAtTile * kAtKRow;    // 17408
constexpr uint32_t kAtVPart = kAtTile * 256;        // 16384
constexpr uint32_t kAtSlot = 2 * kAtKPart + 2 * kAtVPart;  // 67584
constexpr int kAtSlots = 3;
constexpr uint32_t kAtQ = kAtSlots * kAtSlot;       // fp32 [8][128]
constexpr uint32_t kAtP = kAtQ + 8 * 128 * 4;       // fp32 [8][64]
constexpr uint32_t kAtSp = kAtP + 8 * 64 * 4;       // fp32 [4][8][64]
constexpr uint32_t kAtStat = kAtSp + 4 * 8 * 64 * 4;  // fp32 corr[8], linv[8]
constexpr uint32_t kAtEnd = kAtStat + 64;
static_assert(kAtEnd <= kStRingBytes + kStEpiBytes, "attention overlay must fit the ring + epilogue staging");
static_assert(8 * 8 * 128 * 4 <= kAtSlot, "output partials alias tile slot 0");
static_assert(kStSmemBytes <= 232448, "shared memory budget");

enum StackOpKind { kStGemm = 0, kStLayerNorm = 1, kStFsmn = 2, kStAttention = 3 };

struct StackGemmOp {
  int n, k_blocks, kp_a, kp_w, nparts, a_col0, act, taps;
  int tap_row[4];
  int ld_res1, ld_out, f32_col0, out_parts, out_kp, map0, pad0, pad1;
  const float* bias;
  const float* res1;
  float* out_f32;
  __nv_bfloat16* out_split;
  const char* w_base;
  unsigned long long w_bytes;
};
struct StackLnOp {
  const float* x;
  const float* gamma;
  const float* beta;
  float* out_f32;
  __nv_bfloat16* out_split;
  const int32_t* cu_rows;
  int ld_x, dim, ld_f32, nparts, halo;
  float eps;
};
struct StackFsmnOp {
  const float* v;
  const int32_t* cu_rows;
  const float* weight;
  const float* res;
  float* out;
  int ld_v, dim, ksize, ld_res, ld_out;
};
struct StackAttnOp {
  const __nv_bfloat16* q;
  const __nv_bfloat16* k;
  const __nv_bfloat16* v;
  const int32_t* cu_q;
  const int32_t* cu_kv;
  float* out_f32;
  __nv_bfloat16* out_split;
  int ld_q, q_ps, ld_k, ld_v, kv_ps, nparts, heads, ld_out, out_parts, out_kp;
  float scale;
};
struct __align__(16) StackOp {
  int kind, barrier_after, rows, rows_slot, batch, batch_slot, next_gemm, pad;
  union {
    StackGemmOp g;
    StackLnOp ln;
    StackFsmnOp f;
    StackAttnOp a;
  } u;
};
static_assert(sizeof(StackOp) <= kStOpBytes && sizeof(StackOp) % 16 == 0, "op record size");

struct StackDims {
  int v[kStackMaxDims];
};

__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// Device-wide barrier: every thread's global writes of the phase (generic and, through the waits on the bulk groups,
// TMA stores) are ordered before thread 0's release; the waiting side acquires and re-orders the async proxy (the next
// phase reads through TMA / bulk copies what other SMs wrote with plain stores, and the other way round).
__device__ __forceinline__ void grid_barrier(unsigned int* counter, unsigned int& gen) {
  fence_proxy_async_all();
  __syncthreads();
  if (threadIdx.x == 0) {
    ++gen;
    __threadfence();
    atomicAdd(counter, 1u);
    const unsigned int target = gen * gridDim.x;
    unsigned int spins = 0;
    while (ld_acquire_u32(counter) < target) {
      if (++spins == (1u << 23)) {
        printf("fb200: stack kernel grid barrier %u timed out (block %d sees %u of %u)\n", gen, blockIdx.x,
               ld_acquire_u32(counter), target);
        __trap();
      }
    }
    __threadfence();
  } else {
    ++gen;
  }
  __syncthreads();
  fence_proxy_async_all();
}

struct TileCfg {
  int bm, bn, tiles_m, tiles_n;
};
// Tile shape for a [m x ncols x K] product on `grid` CTAs: fewest rounds first, then the smallest per-CTA time estimate
// max(ingest bytes / 41 B/clk, MMA cycles) + a fixed cost per tile.
__device__ __forceinline__ TileCfg pick_tile_cfg(int m, int ncols, int k_iters, int nparts, int grid) {
  TileCfg best = {128, 64, 1, 1};
  long long best_cost = -1;
#pragma unroll
  for (int bi = 0; bi < 4; ++bi) {
    const int bm = 128 - 32 * bi;
#pragma unroll
    for (int ni = 0; ni < 2; ++ni) {
      const int bn = ni == 0 ? 64 : 32;
      const int tm = (m + bm - 1) / bm, tn = (ncols + bn - 1) / bn;
      const int tiles = tm * tn;
      const int rounds = (tiles + grid - 1) / grid;
      const int rows = m < bm ? m : bm;
      const long long ingest = (long long)(((rows + 31) / 32) * 32 + bn) * 128 * nparts * k_iters / 41;
      const long long mma = (long long)k_iters * 4 * (nparts == 1 ? 1 : 3) * (bn / 2);
      const long long cost = (long long)rounds * ((ingest > mma ? ingest : mma) + 1500);
      if (best_cost < 0 || cost < best_cost) {
        best_cost = cost;
        best.bm = bm;
        best.bn = bn;
        best.tiles_m = tm;
        best.tiles_n = tn;
      }
    }
  }
  return best;
}

// ---------------------------------------------------------------- LayerNorm rows ---------
// Same arithmetic as layernorm_kernel<NV> (rowwise.cu): warp per row, row in registers, two-pass variance.
template <int NV>
__device__ __forceinline__ void stack_ln_rows(const StackLnOp& o, int rows, int batch, int warp_global, int warp_stride,
                                              int lane) {
  const int dim = o.dim, nvec = dim >> 2, nparts = o.nparts;
  const int kp = round_up(dim, 64);
  for (int row = warp_global; row < rows; row += warp_stride) {
    const float4* xr = reinterpret_cast<const float4*>(o.x + (size_t)row * o.ld_x);
    float4 v[NV];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = lane + i * 32;
      v[i] = idx < nvec ? __ldcg(xr + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
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
    const float rstd = rsqrtf(warp_sum(q) / (float)dim + o.eps);
    int out_row = row;
    if (o.halo) {
      const int b = find_seq(o.cu_rows, batch, row);
      out_row = row + 1 + b;
      if (o.out_split) {
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        const int nv16 = nparts * kp / 8;
        if (row == __ldg(o.cu_rows + b)) {
          uint4* zr = reinterpret_cast<uint4*>(o.out_split + (size_t)(out_row - 1) * (nparts * kp));
          for (int c = lane; c < nv16; c += 32) zr[c] = z;
        }
        if (row == rows - 1) {
          uint4* zr = reinterpret_cast<uint4*>(o.out_split + (size_t)(out_row + 1) * (nparts * kp));
          for (int c = lane; c < nv16; c += 32) zr[c] = z;
        }
      }
    }
    __nv_bfloat16* srow = o.out_split ? o.out_split + (size_t)out_row * (nparts * kp) : nullptr;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = lane + i * 32;
      if (idx < nvec) {
        const float4 g = __ldg(reinterpret_cast<const float4*>(o.gamma) + idx);
        const float4 b = __ldg(reinterpret_cast<const float4*>(o.beta) + idx);
        float4 y;
        y.x = (v[i].x - mean) * rstd * g.x + b.x;
        y.y = (v[i].y - mean) * rstd * g.y + b.y;
        y.z = (v[i].z - mean) * rstd * g.z + b.z;
        y.w = (v[i].w - mean) * rstd * g.w + b.w;
        if (o.out_f32) reinterpret_cast<float4*>(o.out_f32 + (size_t)row * o.ld_f32)[idx] = y;
        if (srow) {
          const float yy[4] = {y.x, y.y, y.z, y.w};
          store_split4(srow, kp, idx * 4, yy, nparts);
        }
      }
    }
    if (srow) {
      for (int c = dim + lane; c < kp; c += 32)
        for (int pp = 0; pp < nparts; ++pp) srow[pp * kp + c] = __float2bfloat16_rn(0.f);
    }
  }
}

// 8 consecutive bf16 (one LDS.128 / LDG.128) -> fp32
__device__ __forceinline__ void bf16x8_to_f32(const uint4& r, float (&f)[8]) {
  f[0] = __uint_as_float(r.x << 16); f[1] = __uint_as_float(r.x & 0xffff0000u);
  f[2] = __uint_as_float(r.y << 16); f[3] = __uint_as_float(r.y & 0xffff0000u);
  f[4] = __uint_as_float(r.z << 16); f[5] = __uint_as_float(r.z & 0xffff0000u);
  f[6] = __uint_as_float(r.w << 16); f[7] = __uint_as_float(r.w & 0xffff0000u);
}
__device__ __forceinline__ float ld_shared_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void st_shared_f32(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ float4 ld_shared_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint2 ld_shared_u2(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr) : "memory");
  return v;
}

// ---------------------------------------------------------------- the kernel -------------
__global__ void __launch_bounds__(kStackThreads, 1)
stack_kernel(const StackOp* __restrict__ ops, const CUtensorMap* __restrict__ maps, int n_ops, const StackDims dims,
             unsigned int* __restrict__ sync_words) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t epi_base = smem_base + kStRingBytes;
  const uint32_t bar_base = epi_base + kStEpiBytes;
  const uint32_t op_base = bar_base + kStMiscBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kStStages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * kStStages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * kStStages + 2 + a); };
  auto att_bar = [&](int s) { return bar_base + 8u * (2 * kStStages + 4 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kStStages + 4 + kAtSlots);
  const uint32_t bias_sm = bar_base + 256u;  // 64 fp32
  const uint32_t item_sm = bar_base + 512u;  // attention work split: [0] query rows per item, [1] items
  const StackOp* s_op = reinterpret_cast<const StackOp*>(smem_raw + (op_base - smem_u32(smem_raw)));

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int grid = gridDim.x;

  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < kStStages; ++s) {
        mbar_init(full_bar(s), 1);
        mbar_init(empty_bar(s), 1);
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(tfull_bar(a), 1);
        mbar_init(tempty_bar(a), 4);
      }
      for (int s = 0; s < kAtSlots; ++s) mbar_init(att_bar(s), 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(kStTmemCols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  int it = 0;   // k-block stages issued / consumed so far (warps 0 and 1 advance it identically)
  int tl = 0;   // tiles so far (warps 1..5)
  uint32_t att_par = 0;  // bit s: parity of the next phase of attention slot barrier s (all threads advance it alike)
  unsigned int gen = 0;
  int prefetched = -1;  // index of the last GEMM op whose weights were sent towards L2

  for (int oi = 0; oi < n_ops; ++oi) {
    if (tid < (int)(sizeof(StackOp) / 16))
      reinterpret_cast<uint4*>(smem_raw + (op_base - smem_u32(smem_raw)))[tid] = __ldg(reinterpret_cast<const uint4*>(ops + oi) + tid);
    __syncthreads();
    const StackOp& op = *s_op;
    const int rows = op.rows_slot >= 0 ? dims.v[op.rows_slot] : op.rows;
    const int batch = op.batch_slot >= 0 ? dims.v[op.batch_slot] : op.batch;

    // weights of the next GEMM op towards L2 while this op runs (each CTA its share; a hint, nothing waits for it)
    if (op.next_gemm > prefetched && op.next_gemm < n_ops) {
      prefetched = op.next_gemm;
      if (warp == 7) {
        const StackOp* nx = ops + op.next_gemm;
        const char* wb = nx->u.g.w_base;
        const unsigned long long total = nx->u.g.w_bytes;
        const unsigned long long chunk = 4096;
        for (unsigned long long off = ((unsigned long long)blockIdx.x * 32 + lane) * chunk; off < total;
             off += (unsigned long long)grid * 32 * chunk) {
          const unsigned long long left = total - off;
          const uint32_t nbytes = (uint32_t)(left < chunk ? (left & ~15ull) : chunk);
          if (nbytes) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(wb + off), "r"(nbytes) : "memory");
        }
        if (lane == 0) {
          asm volatile("prefetch.tensormap [%0];" ::"l"(maps + nx->u.g.map0) : "memory");
          asm volatile("prefetch.tensormap [%0];" ::"l"(maps + nx->u.g.map0 + 1) : "memory");
          asm volatile("prefetch.tensormap [%0];" ::"l"(maps + nx->u.g.map0 + 2) : "memory");
          asm volatile("prefetch.tensormap [%0];" ::"l"(maps + nx->u.g.map0 + 3) : "memory");
        }
      }
    }

    if (op.kind == kStGemm) {
      // ================================================================ GEMM ============
      const StackGemmOp& g = op.u.g;
      const int m = rows, n = g.n, nparts = g.nparts;
      const int ncols = g.out_split ? g.out_kp : n;
      const int k_iters = g.taps * g.k_blocks;
      const TileCfg cfg = pick_tile_cfg(m, ncols, k_iters, nparts, grid);
      const int bm = cfg.bm, bn = cfg.bn;
      const int total_tiles = cfg.tiles_m * cfg.tiles_n;
      const int my_tiles = total_tiles > (int)blockIdx.x ? (total_tiles - (int)blockIdx.x + grid - 1) / grid : 0;
      const CUtensorMap* map_a = maps + g.map0;
      const CUtensorMap* map_w = maps + g.map0 + 1;
      const CUtensorMap* map_f32 = maps + g.map0 + 2;
      const CUtensorMap* map_split = maps + g.map0 + 3;
      const int nterms = nparts == 1 ? 1 : 3;

      if (warp == 0) {
        if (lane == 0) {  // ===== TMA producer =====
          int i = it;
          const int w_boxes = bn / 32;
          for (int tile = blockIdx.x; tile < total_tiles; tile += grid) {
            const int m0 = (tile / cfg.tiles_n) * bm, n0 = (tile % cfg.tiles_n) * bn;
            const int rows_here = m - m0 < bm ? m - m0 : bm;
            const int a_boxes = (rows_here + 31) / 32;
            for (int tap = 0; tap < g.taps; ++tap) {
              const int a_r = m0 + g.tap_row[tap];
              for (int kb = 0; kb < g.k_blocks; ++kb, ++i) {
                const int s = i % kStStages;
                const uint32_t ph = (uint32_t)(i / kStStages) & 1u;
                mbar_wait(empty_bar(s), ph ^ 1u);
                const uint32_t sa = smem_base + s * kStStageBytes;
                mbar_expect_tx(full_bar(s), (uint32_t)(nparts * (a_boxes + w_boxes)) * 4096u);
                for (int pa = 0; pa < nparts; ++pa)
                  for (int b = 0; b < a_boxes; ++b)
                    tma_load_2d(sa + pa * kStAPart + b * 4096u, map_a, g.a_col0 + pa * g.kp_a + kb * SBK, a_r + b * 32, full_bar(s));
                for (int pb = 0; pb < nparts; ++pb)
                  for (int b = 0; b < w_boxes; ++b)
                    tma_load_2d(sa + 2 * kStAPart + pb * kStBPart + b * 4096u, map_w, pb * g.kp_w + kb * SBK, tap * n + n0 + b * 32,
                                full_bar(s));
              }
            }
          }
        }
        it += my_tiles * k_iters;
        __syncwarp();
      } else if (warp == 1) {
        if (lane == 0) {  // ===== MMA issuer =====
          const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
          int i = it, t = tl;
          for (int tile = blockIdx.x; tile < total_tiles; tile += grid, ++t) {
            const int acc = t & 1;
            mbar_wait(tempty_bar(acc), (((uint32_t)t >> 1) & 1u) ^ 1u);
            tcgen05_fence_after();
            const uint32_t tmem_d = tmem_base + (uint32_t)(acc * 64);
            for (int j = 0; j < k_iters; ++j, ++i) {
              const int s = i % kStStages;
              const uint32_t ph = (uint32_t)(i / kStStages) & 1u;
              mbar_wait(full_bar(s), ph);
              tcgen05_fence_after();
              const uint32_t sa = smem_base + s * kStStageBytes;
              const uint32_t sb = sa + 2 * kStAPart;
              for (int tt = 0; tt < nterms; ++tt) {
                const int ta = (nparts == 2 && tt == 1) ? 1 : 0, tb = (nparts == 2 && tt == 0) ? 1 : 0;
#pragma unroll
                for (int k = 0; k < SBK / 16; ++k) {
                  const uint64_t da = umma_desc_sw128(sa + ta * kStAPart + k * 32);
                  const uint64_t db = umma_desc_sw128(sb + tb * kStBPart + k * 32);
                  umma_bf16(tmem_d, da, db, idesc, (j > 0 || tt > 0 || k > 0) ? 1u : 0u);
                }
              }
              umma_commit(empty_bar(s));
            }
            umma_commit(tfull_bar(acc));
          }
        }
        it += my_tiles * k_iters;
        tl += my_tiles;
        __syncwarp();
      } else if (warp < 6) {
        // ===== epilogue warps 2..5: TMEM lane group = warp % 4, lane = output row =====
        const int lg = warp & 3;
        const uint32_t fbuf = epi_base + (uint32_t)(warp - 2) * kStEpiWarpBytes;
        const uint32_t sbuf = fbuf + 4096u;
        const int f32_col0 = g.f32_col0, out_parts = g.out_parts, out_kp = g.out_kp;
        const float* __restrict__ bias = g.bias;
        const float* res1 = g.res1;
        const bool has_f32 = g.out_f32 != nullptr, has_split = g.out_split != nullptr;
        const float lo = g.act == 1 ? 0.f : -INFINITY;
        const bool swish = g.act == 2;
        const uint32_t sw128 = (uint32_t)(lane & 7), sw64 = (uint32_t)((lane >> 1) & 3);
        const int et = (warp - 2) * 32 + lane;
        int t = tl;
        for (int tile = blockIdx.x; tile < total_tiles; tile += grid, ++t) {
          const int acc = t & 1;
          const int m0 = (tile / cfg.tiles_n) * bm, n0 = (tile % cfg.tiles_n) * bn;
          const int rows_here = m - m0 < bm ? m - m0 : bm;
          if (bias != nullptr) {
            asm volatile("bar.sync 1, 128;" ::: "memory");  // every warp has read the previous tile's values
            if (et < bn) st_shared_f32(bias_sm + (uint32_t)et * 4u, n0 + et < n ? __ldg(bias + n0 + et) : 0.f);
            asm volatile("bar.sync 1, 128;" ::: "memory");
          }
          mbar_wait(tfull_bar(acc), ((uint32_t)t >> 1) & 1u);
          tcgen05_fence_after();
          const bool warp_live = lg * 32 < rows_here;  // this warp's 32 rows belong to the tile
          const int row = m0 + lg * 32 + lane;
          const bool row_ok = row < m;
#pragma unroll 1
          for (int c = 0; c < bn; c += 32) {
            const int col0 = n0 + c;
            const bool live = warp_live && col0 < ncols;
            const bool full = col0 + 32 <= n;
            float y[32];
            if (live) {
              uint32_t r[32];
              tmem_ld32(tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(acc * 64 + c), r);
#pragma unroll
              for (int j = 0; j < 32; ++j) y[j] = __uint_as_float(r[j]);
            }
            if (c + 32 >= bn) {  // accumulator fully read: hand the buffer back to the MMA warp
              tcgen05_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(tempty_bar(acc));
            }
            if (!live) continue;
            if (bias != nullptr) {
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float4 b4 = ld_shared_f4(bias_sm + (uint32_t)(c + 4 * j) * 4u);
                y[4 * j] += b4.x; y[4 * j + 1] += b4.y; y[4 * j + 2] += b4.z; y[4 * j + 3] += b4.w;
              }
            }
#pragma unroll
            for (int j = 0; j < 32; ++j) y[j] = fmaxf(y[j], lo);
            if (swish) {
#pragma unroll
              for (int j = 0; j < 32; ++j) y[j] = __fdividef(y[j], 1.f + __expf(-y[j]));
            }
            if (res1 != nullptr && row_ok) {
              const float* rp = res1 + (size_t)row * g.ld_res1 + col0;
              if (full && (g.ld_res1 & 3) == 0) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  const float4 q = __ldcg(reinterpret_cast<const float4*>(rp) + j);
                  y[4 * j] += q.x; y[4 * j + 1] += q.y; y[4 * j + 2] += q.z; y[4 * j + 3] += q.w;
                }
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (col0 + j < n) y[j] += __ldcg(rp + j);
              }
            }
            if (!full) {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (col0 + j >= n) y[j] = 0.f;
            }
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncwarp();
            const bool do_f32 = has_f32 && col0 >= f32_col0;
            if (do_f32) {
#pragma unroll
              for (int j = 0; j < 8; ++j)
                asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(fbuf + (uint32_t)lane * 128u + ((uint32_t)j ^ sw128) * 16u),
                             "f"(y[4 * j]), "f"(y[4 * j + 1]), "f"(y[4 * j + 2]), "f"(y[4 * j + 3])
                             : "memory");
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
                      y[j] -= __low2float(h);
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
            if (lane == 0) {
              const int r0 = m0 + lg * 32;
              if (do_f32) tma_store_2d(map_f32, fbuf, col0 - f32_col0, r0);
              if (has_split) {
                tma_store_2d(map_split, sbuf, col0, r0);
                if (out_parts > 1) tma_store_2d(map_split, sbuf + 2048u, out_kp + col0, r0);
              }
              asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
          }
        }
        tl += my_tiles;
        if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // this warp's stores have landed
        __syncwarp();
      }
    } else if (op.kind == kStLayerNorm) {
      // ================================================================ LayerNorm =======
      const StackLnOp& o = op.u.ln;
      const int wg = blockIdx.x + grid * warp, ws = grid * (kStackThreads / 32);
      if (o.dim <= 512) stack_ln_rows<4>(o, rows, batch, wg, ws, lane);
      else if (o.dim <= 640) stack_ln_rows<5>(o, rows, batch, wg, ws, lane);
      else stack_ln_rows<16>(o, rows, batch, wg, ws, lane);
    } else if (op.kind == kStFsmn) {
      // ================================================================ FSMN ============
      // out[r, c] = (v[r, c] (+ res[r, c])) + sum_j w[c, j] v[r + j - left, c]  (taps inside the utterance of r, in tap
      // order, one fma chain: fsmn_stream_kernel's sums)
      const StackFsmnOp& o = op.u.f;
      const int nvec = o.dim >> 2, ks = o.ksize, left = (ks - 1) / 2;
      const int total = rows * nvec;
      const int wg = blockIdx.x + grid * warp;
      for (int idx = wg * 32 + lane; idx < total; idx += grid * kStackThreads) {
        const int r = idx / nvec, c = (idx - r * nvec) * 4;
        const int b = find_seq(o.cu_rows, batch, r);
        const int beg = __ldg(o.cu_rows + b), end = __ldg(o.cu_rows + b + 1);
        float4 conv = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int j = 0; j < ks; ++j) {
          const int rr = r - left + j;
          if (rr >= beg && rr < end) {
            const float4 x = __ldcg(reinterpret_cast<const float4*>(o.v + (size_t)rr * o.ld_v + c));
            conv.x = fmaf(__ldg(o.weight + (size_t)(c + 0) * ks + j), x.x, conv.x);
            conv.y = fmaf(__ldg(o.weight + (size_t)(c + 1) * ks + j), x.y, conv.y);
            conv.z = fmaf(__ldg(o.weight + (size_t)(c + 2) * ks + j), x.z, conv.z);
            conv.w = fmaf(__ldg(o.weight + (size_t)(c + 3) * ks + j), x.w, conv.w);
          }
        }
        float4 acc = __ldcg(reinterpret_cast<const float4*>(o.v + (size_t)r * o.ld_v + c));
        if (o.res) {
          const float4 rr = __ldcg(reinterpret_cast<const float4*>(o.res + (size_t)r * o.ld_res + c));
          acc.x += rr.x; acc.y += rr.y; acc.z += rr.z; acc.w += rr.w;
        }
        acc.x += conv.x; acc.y += conv.y; acc.z += conv.z; acc.w += conv.w;
        *reinterpret_cast<float4*>(o.out + (size_t)r * o.ld_out + c) = acc;
      }
    } else if (op.kind == kStAttention) {
      // ================================================================ attention =======
      // item = (utterance, head, AQ <= 8 consecutive query rows); keys / values of the utterance stream through a ring of
      // 64-key tiles (bulk copies of the bf16 parts as they lie in the split matrices); fp32 streaming softmax.
      const StackAttnOp& o = op.u.a;
      const int heads = o.heads, nparts = o.nparts;
      if (tid == 0) {  // query rows per item: the smallest AQ whose item count fits one round (capped at 8)
        const int total_q = __ldg(o.cu_q + batch);
        int aq = (heads * total_q + grid - 1) / grid;
        aq = aq < 1 ? 1 : (aq > 8 ? 8 : aq);
        int items = 0;
        for (;;) {
          items = 0;
          for (int b = 0; b < batch; ++b) {
            const int len = __ldg(o.cu_q + b + 1) - __ldg(o.cu_q + b);
            const int kv = __ldg(o.cu_kv + b + 1) - __ldg(o.cu_kv + b);
            if (kv > 0) items += ((len + aq - 1) / aq) * heads;
          }
          if (items <= grid || aq >= 8) break;
          ++aq;
        }
        asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(item_sm), "r"(aq), "r"(items) : "memory");
      }
      __syncthreads();
      int aq, items;
      asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(aq), "=r"(items) : "r"(item_sm) : "memory");
      for (int item = blockIdx.x; item < items; item += grid) {
        // item -> (utterance, chunk, head)
        int b = 0, rem = item, q_beg = 0, q_len = 0, kv_beg = 0, kv_len = 0;
        for (; b < batch; ++b) {
          q_beg = __ldg(o.cu_q + b);
          q_len = __ldg(o.cu_q + b + 1) - q_beg;
          kv_beg = __ldg(o.cu_kv + b);
          kv_len = __ldg(o.cu_kv + b + 1) - kv_beg;
          const int nb = kv_len > 0 ? ((q_len + aq - 1) / aq) * heads : 0;
          if (rem < nb) break;
          rem -= nb;
        }
        const int head = rem % heads, q0 = (rem / heads) * aq;
        const int nq = q_len - q0 < aq ? q_len - q0 : aq;
        const int hoff = head * 128;
        const int ntiles = (kv_len + kAtTile - 1) / kAtTile;
        auto issue_tile = [&](int t) {  // warp 0: bulk copies of tile t into slot t % 3
          const int s = t % kAtSlots;
          const int nvalid = kv_len - t * kAtTile < kAtTile ? kv_len - t * kAtTile : kAtTile;
          if (lane == 0) mbar_expect_tx(att_bar(s), (uint32_t)(nvalid * nparts * 2) * 256u);
          __syncwarp();
          const uint32_t slot = smem_base + (uint32_t)s * kAtSlot;
          for (int j = lane; j < nvalid; j += 32) {
            const size_t row = (size_t)(kv_beg + t * kAtTile + j);
            for (int pp = 0; pp < nparts; ++pp) {
              const __nv_bfloat16* ks = o.k + row * o.ld_k + hoff + pp * o.kv_ps;
              const __nv_bfloat16* vs = o.v + row * o.ld_v + hoff + pp * o.kv_ps;
              asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                               slot + (uint32_t)pp * kAtKPart + (uint32_t)j * kAtKRow),
                           "l"(ks), "r"(256u), "r"(att_bar(s))
                           : "memory");
              asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                               slot + 2 * kAtKPart + (uint32_t)pp * kAtVPart + (uint32_t)j * 256u),
                           "l"(vs), "r"(256u), "r"(att_bar(s))
                           : "memory");
            }
          }
        };
        if (warp == 0)
          for (int t = 0; t < ntiles && t < kAtSlots; ++t) issue_tile(t);
        // Q rows, pre-scaled like the reference (q * d_k^-0.5 before the product)
        if (tid < 8 * 32) {
          const int i = tid >> 5, c = (tid & 31) * 4;
          float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
          if (i < nq) {
            const __nv_bfloat16* qp = o.q + (size_t)(q_beg + q0 + i) * o.ld_q + hoff + c;
            for (int pp = 0; pp < nparts; ++pp) {
              const uint2 raw = __ldcg(reinterpret_cast<const uint2*>(qp + (size_t)pp * o.q_ps));
              val.x += __uint_as_float(raw.x << 16); val.y += __uint_as_float(raw.x & 0xffff0000u);
              val.z += __uint_as_float(raw.y << 16); val.w += __uint_as_float(raw.y & 0xffff0000u);
            }
            val.x *= o.scale; val.y *= o.scale; val.z *= o.scale; val.w *= o.scale;
          }
          asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(smem_base + kAtQ + (uint32_t)(i * 128 + c) * 4u), "f"(val.x),
                       "f"(val.y), "f"(val.z), "f"(val.w)
                       : "memory");
        }
        __syncthreads();

        float oacc[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int u = 0; u < 4; ++u) oacc[i][u] = 0.f;
        float m_run = -INFINITY, l_run = 0.f;  // of row `warp` (uniform across the warp)
        const int key = tid & 63, dq = tid >> 6;  // S phase: one key, one quarter of the head dimension
        const int c4 = tid & 31, kq = tid >> 5;   // PV phase: four output columns, one eighth of the tile's keys

        for (int t = 0; t < ntiles; ++t) {
          const int s = t % kAtSlots;
          const int nvalid = kv_len - t * kAtTile < kAtTile ? kv_len - t * kAtTile : kAtTile;
          mbar_wait(att_bar(s), (att_par >> s) & 1u);
          att_par ^= 1u << s;
          const uint32_t slot = smem_base + (uint32_t)s * kAtSlot;
          // ---- partial scores: S[i][key] over dims [dq*32, dq*32+32)
          {
            float sp[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) sp[i] = 0.f;
            if (key < nvalid) {
#pragma unroll
              for (int d8 = 0; d8 < 4; ++d8) {
                const uint32_t koff = (uint32_t)key * kAtKRow + (uint32_t)(dq * 64 + d8 * 16);
                uint4 raw;
                float kf[8];
                ld_shared_v4(slot + koff, raw);
                bf16x8_to_f32(raw, kf);
                if (nparts > 1) {
                  float km[8];
                  ld_shared_v4(slot + kAtKPart + koff, raw);
                  bf16x8_to_f32(raw, km);
#pragma unroll
                  for (int u = 0; u < 8; ++u) kf[u] += km[u];
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                  if (i < aq) {
                    const uint32_t qa = smem_base + kAtQ + (uint32_t)(i * 128 + dq * 32 + d8 * 8) * 4u;
                    const float4 q0v = ld_shared_f4(qa), q1v = ld_shared_f4(qa + 16u);
                    sp[i] = fmaf(q0v.x, kf[0], sp[i]); sp[i] = fmaf(q0v.y, kf[1], sp[i]);
                    sp[i] = fmaf(q0v.z, kf[2], sp[i]); sp[i] = fmaf(q0v.w, kf[3], sp[i]);
                    sp[i] = fmaf(q1v.x, kf[4], sp[i]); sp[i] = fmaf(q1v.y, kf[5], sp[i]);
                    sp[i] = fmaf(q1v.z, kf[6], sp[i]); sp[i] = fmaf(q1v.w, kf[7], sp[i]);
                  }
                }
              }
            }
#pragma unroll
            for (int i = 0; i < 8; ++i)
              if (i < aq) st_shared_f32(smem_base + kAtSp + (uint32_t)((dq * 8 + i) * 64 + key) * 4u, sp[i]);
          }
          __syncthreads();
          // ---- softmax of row `warp` over this tile (lanes: keys lane, lane + 32), running max / sum
          if (warp < aq) {
            float s0 = -INFINITY, s1 = -INFINITY;
            if (lane < nvalid) {
              s0 = 0.f;
#pragma unroll
              for (int d = 0; d < 4; ++d) s0 += ld_shared_f32(smem_base + kAtSp + (uint32_t)((d * 8 + warp) * 64 + lane) * 4u);
            }
            if (lane + 32 < nvalid) {
              s1 = 0.f;
#pragma unroll
              for (int d = 0; d < 4; ++d) s1 += ld_shared_f32(smem_base + kAtSp + (uint32_t)((d * 8 + warp) * 64 + lane + 32) * 4u);
            }
            const float mx = warp_max(fmaxf(s0, s1));
            const float m_new = fmaxf(m_run, mx);
            const float corr = expf(m_run - m_new);  // exp(-inf) = 0 on the first tile
            const float p0 = lane < nvalid ? expf(s0 - m_new) : 0.f;
            const float p1 = lane + 32 < nvalid ? expf(s1 - m_new) : 0.f;
            const float sum = warp_sum(p0 + p1);
            l_run = l_run * corr + sum;
            m_run = m_new;
            st_shared_f32(smem_base + kAtP + (uint32_t)(warp * 64 + lane) * 4u, p0);
            st_shared_f32(smem_base + kAtP + (uint32_t)(warp * 64 + lane + 32) * 4u, p1);
            if (lane == 0) st_shared_f32(smem_base + kAtStat + (uint32_t)warp * 4u, corr);
          }
          __syncthreads();
          // ---- O[i][c4*4 .. +3] (partial over this thread's eighth of the keys) = O * corr + P . V
          {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              if (i < aq) {
                const float corr = ld_shared_f32(smem_base + kAtStat + (uint32_t)i * 4u);
#pragma unroll
                for (int u = 0; u < 4; ++u) oacc[i][u] *= corr;
              }
            }
#pragma unroll 2
            for (int kk = 0; kk < 8; ++kk) {
              const int ky = kq * 8 + kk;
              if (ky < nvalid) {
                const uint32_t voff = slot + 2 * kAtKPart + (uint32_t)ky * 256u + (uint32_t)c4 * 8u;
                uint2 raw = ld_shared_u2(voff);
                float v0 = __uint_as_float(raw.x << 16), v1 = __uint_as_float(raw.x & 0xffff0000u);
                float v2 = __uint_as_float(raw.y << 16), v3 = __uint_as_float(raw.y & 0xffff0000u);
                if (nparts > 1) {
                  raw = ld_shared_u2(voff + kAtVPart);
                  v0 += __uint_as_float(raw.x << 16); v1 += __uint_as_float(raw.x & 0xffff0000u);
                  v2 += __uint_as_float(raw.y << 16); v3 += __uint_as_float(raw.y & 0xffff0000u);
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                  if (i < aq) {
                    const float p = ld_shared_f32(smem_base + kAtP + (uint32_t)(i * 64 + ky) * 4u);
                    oacc[i][0] = fmaf(p, v0, oacc[i][0]); oacc[i][1] = fmaf(p, v1, oacc[i][1]);
                    oacc[i][2] = fmaf(p, v2, oacc[i][2]); oacc[i][3] = fmaf(p, v3, oacc[i][3]);
                  }
                }
              }
            }
          }
          __syncthreads();  // tile consumed: its slot, the score partials and P may be overwritten
          if (warp == 0 && t + kAtSlots < ntiles) issue_tile(t + kAtSlots);
        }
        // ---- sum the eight key-eighths (fixed order), normalise, store
        if (warp < aq && lane == 0) st_shared_f32(smem_base + kAtStat + 32u + (uint32_t)warp * 4u, 1.f / l_run);
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (i < aq)
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(smem_base + (uint32_t)(((kq * 8 + i) * 128) + c4 * 4) * 4u),
                         "f"(oacc[i][0]), "f"(oacc[i][1]), "f"(oacc[i][2]), "f"(oacc[i][3])
                         : "memory");
        __syncthreads();
        {
          const int i = tid >> 5;  // row i, columns c4*4..+3
          if (i < nq) {
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const float4 pv = ld_shared_f4(smem_base + (uint32_t)(((e * 8 + i) * 128) + c4 * 4) * 4u);
              acc.x += pv.x; acc.y += pv.y; acc.z += pv.z; acc.w += pv.w;
            }
            const float inv = ld_shared_f32(smem_base + kAtStat + 32u + (uint32_t)i * 4u);
            const float yy[4] = {acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv};
            const size_t row = (size_t)(q_beg + q0 + i);
            const int col = hoff + c4 * 4;
            if (o.out_f32) *reinterpret_cast<float4*>(o.out_f32 + row * o.ld_out + col) = make_float4(yy[0], yy[1], yy[2], yy[3]);
            if (o.out_split) store_split4(o.out_split + row * (size_t)(o.out_parts * o.out_kp), o.out_kp, col, yy, o.out_parts);
          }
        }
        __syncthreads();  // the partials (tile slot 0) and the statistics are free for the next item
      }
    }

    // ---- end of the op
    if (op.barrier_after)
      grid_barrier(sync_words, gen);
    else
      __syncthreads();
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kStTmemCols));
  }
  if (tid == 0) {  // the last CTA out re-arms the counters for the next run of the program
    __threadfence();
    const unsigned int old = atomicAdd(sync_words + 1, 1u);
    if (old == gridDim.x - 1) {
      sync_words[0] = 0u;
      sync_words[1] = 0u;
      __threadfence();
    }
  }
}

// ================================================================ host: recorder =========
struct ByteRange {
  const char* lo;
  const char* hi;
};
struct StackProgram {
  std::vector<StackOp> ops;
  std::vector<CUtensorMap> maps;
  std::vector<ByteRange> reads, writes;  // since the last barrier
  int dim_values[kStackMaxDims];
  int ndims = 0;
  int barriers = 0;
  bool open = true, failed = false;
  StackOp* d_ops = nullptr;
  CUtensorMap* d_maps = nullptr;
  unsigned int* d_sync = nullptr;
  int grid = 0;
};
static thread_local StackProgram* g_rec = nullptr;

bool stack_recording() { return g_rec != nullptr; }

static bool overlaps(const std::vector<ByteRange>& v, const ByteRange& r) {
  for (const ByteRange& x : v)
    if (r.lo < x.hi && x.lo < r.hi) return true;
  return false;
}
static ByteRange range_of(const void* p, size_t bytes) {
  ByteRange r;
  r.lo = (const char*)p;
  r.hi = r.lo + bytes;
  return r;
}
static int slot_of(const StackProgram& pr, int value) {
  for (int s = 0; s < pr.ndims; ++s)
    if (pr.dim_values[s] == value) return s;
  return -1;
}
// Appends `op`: a barrier goes between it and its predecessor iff it touches what an op since the last barrier wrote, or
// writes what one of them read.
static void append_op(StackProgram& pr, StackOp op, const std::vector<ByteRange>& rd, const std::vector<ByteRange>& wr) {
  bool dep = false;
  for (const ByteRange& r : rd) dep = dep || overlaps(pr.writes, r);
  for (const ByteRange& w : wr) dep = dep || overlaps(pr.writes, w) || overlaps(pr.reads, w);
  if (!pr.ops.empty()) {
    pr.ops.back().barrier_after = dep ? 1 : 0;
    if (dep) {
      ++pr.barriers;
      pr.reads.clear();
      pr.writes.clear();
    }
  }
  for (const ByteRange& r : rd)
    if (r.lo != nullptr && r.hi > r.lo) pr.reads.push_back(r);
  for (const ByteRange& w : wr)
    if (w.lo != nullptr && w.hi > w.lo) pr.writes.push_back(w);
  op.barrier_after = 0;
  op.next_gemm = 1 << 30;
  pr.ops.push_back(op);
}

int stack_record_gemm(const void* a, int a_rows, int lda, int a_parts, const void* w, int w_rows, int ldw, const GemmParams& p) {
  StackProgram& pr = *g_rec;
  StackOp op = {};
  op.kind = kStGemm;
  op.rows = p.m;
  op.rows_slot = slot_of(pr, p.m);
  op.batch = 0;
  op.batch_slot = -1;
  StackGemmOp& g = op.u.g;
  g.n = p.n;
  g.k_blocks = p.k_blocks;
  g.kp_a = p.kp_a;
  g.kp_w = p.kp_w;
  g.nparts = a_parts;
  g.a_col0 = p.a_col0;
  g.act = p.act;
  g.taps = p.taps;
  for (int j = 0; j < 4; ++j) g.tap_row[j] = j < p.taps ? p.tap_row[j] : 0;
  g.ld_res1 = p.ld_res1;
  g.ld_out = p.ld_out;
  g.f32_col0 = p.f32_col0;
  g.out_parts = p.out_parts;
  g.out_kp = p.out_kp;
  g.map0 = (int)pr.maps.size();
  g.bias = p.bias;
  g.res1 = p.res1;
  g.out_f32 = p.out_f32;
  g.out_split = p.out_split;
  g.w_base = (const char*)w;
  g.w_bytes = (unsigned long long)w_rows * ldw * 2ull;
  // operand maps in 32-row boxes (the kernel picks the tile height and width per run); output boxes as in gemm_host.inc
  CUtensorMap ma, mw, mf, ms;
  int rc = make_tma_map(&ma, a, 2, (uint64_t)a_rows, (uint64_t)lda, (uint64_t)lda, 64, 32, 128);
  if (rc) return rc;
  rc = make_tma_map(&mw, w, 2, (uint64_t)w_rows, (uint64_t)ldw, (uint64_t)ldw, 64, 32, 128);
  if (rc) return rc;
  mf = ma;
  ms = ma;
  if (p.out_f32) {
    rc = make_tma_map(&mf, p.out_f32, 4, (uint64_t)p.m, (uint64_t)(p.n - p.f32_col0), (uint64_t)p.ld_out, 32, 32, 128);
    if (rc) return rc;
  }
  if (p.out_split) {
    rc = make_tma_map(&ms, p.out_split, 2, (uint64_t)p.m, (uint64_t)p.out_parts * p.out_kp, (uint64_t)p.out_parts * p.out_kp, 32,
                      32, 64);
    if (rc) return rc;
  }
  pr.maps.push_back(ma);
  pr.maps.push_back(mw);
  pr.maps.push_back(mf);
  pr.maps.push_back(ms);
  std::vector<ByteRange> rd, wr;
  rd.push_back(range_of(a, (size_t)a_rows * lda * 2));
  if (p.res1) rd.push_back(range_of(p.res1, (size_t)p.m * p.ld_res1 * 4));
  if (p.out_f32) wr.push_back(range_of(p.out_f32, (size_t)p.m * p.ld_out * 4));
  if (p.out_split) wr.push_back(range_of(p.out_split, (size_t)p.m * p.out_parts * p.out_kp * 2));
  append_op(pr, op, rd, wr);
  return FB200_OK;
}

int stack_record_layernorm(const float* x, int ld_x, int rows, int dim, const float* gamma, const float* beta, float eps,
                           float* out_f32, int ld_f32, void* out_split, int nparts, int halo, const int32_t* cu_rows,
                           int batch) {
  StackProgram& pr = *g_rec;
  StackOp op = {};
  op.kind = kStLayerNorm;
  op.rows = rows;
  op.rows_slot = slot_of(pr, rows);
  op.batch = batch;
  op.batch_slot = halo ? slot_of(pr, batch) : -1;
  StackLnOp& o = op.u.ln;
  o.x = x;
  o.gamma = gamma;
  o.beta = beta;
  o.out_f32 = out_f32;
  o.out_split = (__nv_bfloat16*)out_split;
  o.cu_rows = cu_rows;
  o.ld_x = ld_x;
  o.dim = dim;
  o.ld_f32 = ld_f32;
  o.nparts = nparts;
  o.halo = halo;
  o.eps = eps;
  const int kp = round_up(dim, 64);
  std::vector<ByteRange> rd, wr;
  rd.push_back(range_of(x, (size_t)rows * ld_x * 4));
  if (out_f32) wr.push_back(range_of(out_f32, (size_t)rows * ld_f32 * 4));
  if (out_split) wr.push_back(range_of(out_split, (size_t)(rows + (halo ? batch + 1 : 0)) * nparts * kp * 2));
  append_op(pr, op, rd, wr);
  return FB200_OK;
}

int stack_record_fsmn(const float* v, int ld_v, const int32_t* cu_rows, int batch, int total_rows, int dim,
                      const float* weight, int ksize, const float* residual, int ld_res, float* out, int ld_out) {
  StackProgram& pr = *g_rec;
  StackOp op = {};
  op.kind = kStFsmn;
  op.rows = total_rows;
  op.rows_slot = slot_of(pr, total_rows);
  op.batch = batch;
  op.batch_slot = slot_of(pr, batch);
  StackFsmnOp& o = op.u.f;
  o.v = v;
  o.cu_rows = cu_rows;
  o.weight = weight;
  o.res = residual;
  o.out = out;
  o.ld_v = ld_v;
  o.dim = dim;
  o.ksize = ksize;
  o.ld_res = ld_res;
  o.ld_out = ld_out;
  std::vector<ByteRange> rd, wr;
  rd.push_back(range_of(v, (size_t)total_rows * ld_v * 4));
  if (residual) rd.push_back(range_of(residual, (size_t)total_rows * ld_res * 4));
  wr.push_back(range_of(out, (size_t)total_rows * ld_out * 4));
  append_op(pr, op, rd, wr);
  return FB200_OK;
}

int stack_record_attention(const void* q, int ld_q, int q_col0, int q_part_stride, const void* k, int ld_k, int k_col0,
                           const void* v, int ld_v, int v_col0, int kv_part_stride, int nparts, const int32_t* cu_q,
                           const int32_t* cu_kv, int batch, int q_rows, int kv_rows, int heads, int head_dim, float scale,
                           float* out_f32, int ld_out, void* out_split, int out_parts) {
  StackProgram& pr = *g_rec;
  FB_CHECK_ARG(head_dim == 128, "attention (stack): head_dim %d unsupported", head_dim);
  FB_CHECK_ARG((ld_k * 2) % 16 == 0 && (ld_v * 2) % 16 == 0 && k_col0 % 8 == 0 && v_col0 % 8 == 0 && kv_part_stride % 8 == 0,
               "attention (stack): key / value rows must be 16-byte aligned");
  StackOp op = {};
  op.kind = kStAttention;
  op.rows = q_rows;
  op.rows_slot = slot_of(pr, q_rows);
  op.batch = batch;
  op.batch_slot = slot_of(pr, batch);
  StackAttnOp& o = op.u.a;
  o.q = (const __nv_bfloat16*)q + q_col0;
  o.k = (const __nv_bfloat16*)k + k_col0;
  o.v = (const __nv_bfloat16*)v + v_col0;
  o.cu_q = cu_q;
  o.cu_kv = cu_kv;
  o.out_f32 = out_f32;
  o.out_split = (__nv_bfloat16*)out_split;
  o.ld_q = ld_q;
  o.q_ps = q_part_stride;
  o.ld_k = ld_k;
  o.ld_v = ld_v;
  o.kv_ps = kv_part_stride;
  o.nparts = nparts;
  o.heads = heads;
  o.ld_out = ld_out;
  o.out_parts = out_split ? out_parts : 0;
  o.out_kp = round_up(heads * head_dim, 64);
  o.scale = scale;
  std::vector<ByteRange> rd, wr;
  rd.push_back(range_of(q, (size_t)q_rows * ld_q * 2));
  rd.push_back(range_of(k, (size_t)kv_rows * ld_k * 2));
  rd.push_back(range_of(v, (size_t)kv_rows * ld_v * 2));
  if (out_f32) wr.push_back(range_of(out_f32, (size_t)q_rows * ld_out * 4));
  if (out_split) wr.push_back(range_of(out_split, (size_t)q_rows * out_parts * o.out_kp * 2));
  append_op(pr, op, rd, wr);
  return FB200_OK;
}

}  // namespace fb200

using namespace fb200;

extern "C" __attribute__((visibility("default"))) int fb200_stack_begin(void** handle, const int32_t* dim_values, int ndims) {
  FB_CHECK_ARG(handle != nullptr, "stack_begin: null handle");
  FB_CHECK_ARG(g_rec == nullptr, "stack_begin: a recording is already open on this thread");
  FB_CHECK_ARG(ndims >= 0 && ndims <= kStackMaxDims && (ndims == 0 || dim_values != nullptr), "stack_begin: ndims %d (0..%d)", ndims,
               kStackMaxDims);
  StackProgram* pr = new StackProgram();
  pr->ndims = ndims;
  for (int i = 0; i < ndims; ++i) {
    pr->dim_values[i] = dim_values[i];
    for (int j = 0; j < i; ++j)
      if (dim_values[j] == dim_values[i]) {
        delete pr;
        set_error("stack_begin: run-time dimensions %d and %d share the recording value %d", j, i, dim_values[i]);
        return FB200_EINVAL;
      }
  }
  g_rec = pr;
  *handle = pr;
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_stack_end(void* handle) {
  StackProgram* pr = (StackProgram*)handle;
  FB_CHECK_ARG(pr != nullptr && pr == g_rec && pr->open, "stack_end: not the recording open on this thread");
  g_rec = nullptr;
  pr->open = false;
  FB_CHECK_ARG(!pr->ops.empty(), "stack_end: empty program");
  // next_gemm: index of the first GEMM op after each op (weights prefetched one op ahead)
  int next = 1 << 30;
  for (int i = (int)pr->ops.size() - 1; i >= 0; --i) {
    pr->ops[i].next_gemm = next;
    if (pr->ops[i].kind == kStGemm) next = i;
  }
  pr->ops.back().barrier_after = 0;
  static int grid = 0;
  if (grid == 0) {
    cudaError_t e = cudaFuncSetAttribute(stack_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kStSmemBytes);
    int per_sm = 0;
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, stack_kernel, kStackThreads, kStSmemBytes);
    if (e != cudaSuccess || per_sm < 1) {
      set_error("stack_end: the stack kernel does not fit an SM: %s", cudaGetErrorString(e));
      pr->failed = true;
      return FB200_ECUDA;
    }
    grid = sm_count();
  }
  pr->grid = grid;
  cudaError_t e = cudaMalloc(&pr->d_ops, pr->ops.size() * sizeof(StackOp));
  if (e == cudaSuccess) e = cudaMalloc(&pr->d_maps, (pr->maps.size() + 1) * sizeof(CUtensorMap));
  if (e == cudaSuccess) e = cudaMalloc(&pr->d_sync, 256);
  if (e == cudaSuccess) e = cudaMemcpy(pr->d_ops, pr->ops.data(), pr->ops.size() * sizeof(StackOp), cudaMemcpyHostToDevice);
  if (e == cudaSuccess && !pr->maps.empty())
    e = cudaMemcpy(pr->d_maps, pr->maps.data(), pr->maps.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemset(pr->d_sync, 0, 256);
  if (e != cudaSuccess) {
    set_error("stack_end: device upload failed: %s", cudaGetErrorString(e));
    pr->failed = true;
    return FB200_ECUDA;
  }
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_stack_info(void* handle, int32_t* n_ops, int32_t* n_barriers) {
  StackProgram* pr = (StackProgram*)handle;
  FB_CHECK_ARG(pr != nullptr, "stack_info: null handle");
  if (n_ops) *n_ops = (int32_t)pr->ops.size();
  if (n_barriers) *n_barriers = pr->barriers;
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_stack_run(void* handle, const int32_t* dims, int ndims, void* stream) {
  StackProgram* pr = (StackProgram*)handle;
  FB_CHECK_ARG(pr != nullptr && !pr->open && !pr->failed && pr->d_ops != nullptr, "stack_run: program not finished");
  FB_CHECK_ARG(ndims == pr->ndims, "stack_run: %d run-time dimensions given, the program has %d", ndims, pr->ndims);
  StackDims d = {};
  for (int i = 0; i < ndims; ++i) {
    FB_CHECK_ARG(dims[i] >= 0 && dims[i] <= pr->dim_values[i], "stack_run: dimension %d = %d exceeds the recorded capacity %d", i, dims[i],
                 pr->dim_values[i]);
    d.v[i] = dims[i];
  }
  const StackOp* ops = pr->d_ops;
  const CUtensorMap* maps = pr->d_maps;
  int n_ops = (int)pr->ops.size();
  unsigned int* sync_words = pr->d_sync;
  void* args[] = {(void*)&ops, (void*)&maps, (void*)&n_ops, (void*)&d, (void*)&sync_words};
  const cudaError_t e = cudaLaunchCooperativeKernel((const void*)stack_kernel, dim3((unsigned)pr->grid), dim3(kStackThreads), args,
                                                    kStSmemBytes, (cudaStream_t)stream);
  if (e != cudaSuccess) {
    set_error("stack_run: launch failed: %s", cudaGetErrorString(e));
    return FB200_ECUDA;
  }
  count_launch();
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_stack_destroy(void* handle) {
  StackProgram* pr = (StackProgram*)handle;
  if (pr == nullptr) return FB200_OK;
  if (g_rec == pr) g_rec = nullptr;
  if (pr->d_ops) cudaFree(pr->d_ops);
  if (pr->d_maps) cudaFree(pr->d_maps);
  if (pr->d_sync) cudaFree(pr->d_sync);
  delete pr;
  return FB200_OK;
}
