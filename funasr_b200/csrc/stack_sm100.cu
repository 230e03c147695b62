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
// Arithmetic: GEMM (unsplit), LayerNorm and FSMN ops compute bit for bit what the per-op kernels compute (same bf16 split
// operands, same term order per k-block, same epilogue, same two-pass LayerNorm, same tap order) -- the tile shapes
// differ, the sums do not.  A product whose only consumer is the LayerNorm op behind it (out-projection, FFN w2) is cut
// into K ranges, summed in a fixed order by that LayerNorm op (last-bit differences to the per-op path).  Attention runs
// on the legacy tensor-core path (mma.sync m16n8k16, bf16 hi / mid operands, the same hi.mid + mid.hi + hi.hi products
// as the tcgen05 kernel) over (utterance, head, <= 16 query rows) items: at this size spreading a few query rows over
// every SM beats eight 128-row tcgen05 items.
//
// What bounds a phase here (all measured, tools/stack_bench.py, tools/micro/*.cu, profiles/r02_request_stack.md):
//   * a TMA instruction occupies its issuing warp for ~180-240 cycles whatever it moves -> operands are fetched as 4-d
//     boxes (k-blocks x parts per instruction) by two producer warps;
//   * tcgen05.mma with a <= 64-column tile costs ~41-55 cycles (A is re-read from shared memory), so the per-CTA time of
//     a product is set by the NUMBER of MMAs (k-blocks x 4 x terms): only a K split shortens it;
//   * TMA / MMA issue must sit behind elect.sync -- behind `lane == 0` the compiler wraps every UTMALDG / UTCHMMA in an
//     ELECT + 7 x R2UR loop (~100 cycles apiece);
//   * a device-wide barrier costs ~1.25 us (release, L2 atomic, acquire poll) + the skew between CTAs.
//
// Small-M tiling: bm (32..128 rows), bn (32 / 64 columns) and the K split are picked on the device per op from the
// ACTUAL row count (by an otherwise idle warp, one op ahead); programs are recorded once for a row CAPACITY and run with
// the request's row counts (`dims`).  The next GEMM's weights and the next op's parameter vectors are prefetched into L2
// (cp.async.bulk.prefetch.L2) while the current op runs.
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
// The operand ring is cut per op into as many stages as fit (a stage = all parts of a [bm x 64] A block and a [bn x 64]
// W block: 12 KB..48 KB): a TMA load takes ~2.5 us to land here, so what sets a small-tile CTA's rate is the BYTES IN
// FLIGHT, not the stage count -- 4 stages of 16 KB measured 23 GB/s per SM.
constexpr int kStMaxStages = 16;
constexpr uint32_t kStRingBytes = 196608;
constexpr uint32_t kStEpiWarpBytes = 8192;
constexpr uint32_t kStEpiBytes = 4 * kStEpiWarpBytes;
constexpr uint32_t kStMiscBytes = 1024;      // barriers, tensor-memory slot (0..511), bias tile (512..767), attention split (768..)
constexpr uint32_t kStOpBytes = 256;
constexpr uint32_t kStSmemBytes = kStRingBytes + kStEpiBytes + kStMiscBytes + kStOpBytes + 1024;
constexpr int kStTmemCols = 128;             // two accumulator buffers of 64 columns
constexpr int kStackMaxDims = 8;

// attention overlay: the ring holds three 64-key tiles (K hi | K mid | V hi | V mid, each two [64 rows x 128 B] TMA boxes
// in the 128-byte swizzle, which also makes the per-key 16-byte reads of the score loop conflict-free); the epilogue
// staging area holds the query rows, the probabilities, the partial scores and the row statistics
constexpr int kAtTile = 64;                         // keys per tile
constexpr uint32_t kAtBox = 64 * 128;               // one [64 x 64] bf16 box
constexpr uint32_t kAtPart = 2 * kAtBox;            // one part of K or V of a tile: dims 0..63 | 64..127
constexpr uint32_t kAtSlot = 4 * kAtPart;           // 65536
constexpr int kAtSlots = 3;
constexpr int kAtRows = 16;                         // query rows per item at most (one m16 MMA row block)
constexpr uint32_t kAtQ = kAtSlots * kAtSlot;       // bf16 [2 parts][16 rows][128 dims], 16-byte chunks XOR-swizzled by the row
constexpr uint32_t kAtP = kAtQ + 2 * 16 * 256;      // bf16 [2 parts][16 rows][64 keys], swizzled alike
constexpr uint32_t kAtStat = kAtP + 2 * 16 * 128;   // fp32 [8 warps][16 rows]: per-warp row maxima, later row sums
constexpr uint32_t kAtEnd = kAtStat + 8 * 16 * 4;
static_assert(kAtSlots * kAtSlot <= kStRingBytes, "attention tiles live in the ring");
static_assert(kAtEnd <= kStRingBytes + kStEpiBytes, "attention overlay must fit the ring + epilogue staging");
static_assert(kStSmemBytes <= 232448, "shared memory budget");

enum StackOpKind { kStGemm = 0, kStLayerNorm = 1, kStFsmn = 2, kStAttention = 3 };

struct StackGemmOp {
  int n, k_blocks, kp_a, kp_w, nparts, a_col0, act, taps;
  int tap_row[4];
  int ld_res1, ld_out, f32_col0, out_parts, out_kp, map0;
  // split-K (set by the recorder when a LayerNorm op consumes this product): up to ksplit_max K ranges per output tile, each
  // range writes its raw fp32 accumulators to rows [ks * part_rows, ...) of the scratch matrix behind map0 / out_f32; bias
  // and residual move into the LayerNorm op
  int ksplit_max, part_rows;
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
  // fused reduction of a split-K product (part != nullptr): x = sum over the K ranges of part[ks] (+ bias) (+ res), stored
  // to `x` (it stays the residual stream) before it is normalised; g_* restate the product so that this op derives the
  // same ksplit as the GEMM op did
  const float* part;
  const float* p_bias;
  const float* p_res;
  long long part_stride;
  int ld_part, ld_res, g_ncols, g_kblocks, g_nparts, ksplit_max;
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
  int map0, k_col0, v_col0;  // maps[map0] / maps[map0 + 1]: [64 x 64] boxes over the whole key / value matrices
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
__device__ __forceinline__ void grid_barrier(unsigned int* counter, unsigned int& gen, int flags) {
  if (!(flags & 4)) fence_proxy_async_all();
  __syncthreads();
  ++gen;
  if (threadIdx.x == 0) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(counter), "r"(1u) : "memory");
    const unsigned int target = gen * gridDim.x;
    unsigned int spins = 0;
    while (ld_acquire_u32(counter) < target) {
      if (++spins == (1u << 23)) {
        printf("fb200: stack kernel grid barrier %u timed out (block %d sees %u of %u)\n", gen, blockIdx.x,
               ld_acquire_u32(counter), target);
        __trap();
      }
    }
  }
  __syncthreads();
  fence_proxy_async_all();
}

// Tile shapes of a GEMM op: [bm x bn] output tile, `kbs` k-blocks per TMA instruction / ring stage (a stage holds
// <= 64 KB so that the ring always has >= 3 of them), and which of the op's pre-encoded operand maps has that box.
// Map order of a GEMM op (from map0): fp32 out, split out, A maps (bm, kbs) = (32,4) (32,2) (64,2) (96,2) (96,1) (128,1),
// W maps (bn, kbs) = (32,4) (32,2) (32,1) (64,2) (64,1).
struct StackTile {
  int bm, bn, kbs, a_map, w_map;
};
constexpr int kStTiles = 8;
__host__ __device__ constexpr StackTile stack_tile(int i) {
  return i == 0 ? StackTile{128, 64, 1, 5, 4} : i == 1 ? StackTile{128, 32, 1, 5, 2} : i == 2 ? StackTile{96, 64, 1, 4, 4}
       : i == 3 ? StackTile{96, 32, 2, 3, 1}  : i == 4 ? StackTile{64, 64, 2, 2, 3}  : i == 5 ? StackTile{64, 32, 2, 2, 1}
       : i == 6 ? StackTile{32, 64, 2, 1, 3}  : StackTile{32, 32, 4, 0, 0};
}
constexpr int kStMapsPerGemm = 13, kStAMap0 = 2, kStWMap0 = 8;
static const int kHostAMaps[6][2] = {{32, 4}, {32, 2}, {64, 2}, {96, 2}, {96, 1}, {128, 1}};
static const int kHostWMaps[5][2] = {{32, 4}, {32, 2}, {32, 1}, {64, 2}, {64, 1}};

struct TileCfg {
  int idx, tiles_m, tiles_n, ksplit;
};
// Tile shape (and K split) for a [m x ncols x K] product on `grid` CTAs: the smallest estimate of
//   rounds * (max(ingest bytes / 64 B/clk, MMA issue) + fixed cost per unit)
// where one MMA costs its issuing thread max(40, bn / 2) cycles: with <= 64-column tiles the per-CTA time is set by the
// NUMBER of MMAs (k-blocks x 4 K-slices x terms), which only a K split shortens.
__device__ __noinline__ TileCfg pick_tile_cfg(int m, int ncols, int k_blocks, int nparts, int grid, int ksplit_max) {
  TileCfg best = {0, 1, 1, 1};
  int best_cost = 0x7fffffff;
  const int terms = nparts == 1 ? 1 : 3;
#pragma unroll
  for (int pass = 0; pass < 2; ++pass) {  // unsplit, then the full split (everything below is division by constants)
    const int ks = pass == 0 ? 1 : ksplit_max;
    if (pass == 1 && (ksplit_max <= 1 || (k_blocks & (ksplit_max - 1)) != 0)) break;
    const int kb_unit = pass == 0 ? k_blocks : k_blocks / ksplit_max;
#pragma unroll
    for (int ci = 0; ci < kStTiles; ++ci) {
      constexpr StackTile t0 = stack_tile(0);
      (void)t0;
      const StackTile t = stack_tile(ci);
      if ((kb_unit & (t.kbs - 1)) != 0) continue;
      const int tm = (m + t.bm - 1) / t.bm, tn = (ncols + t.bn - 1) / t.bn;
      const int units = tm * tn * ks;
      const int rounds = units <= grid ? 1 : (units <= 2 * grid ? 2 : (units + grid - 1) / grid);
      const int ingest = (t.bm + t.bn) * nparts * kb_unit * 2;  // cycles: 128 B per row and k-block at ~64 B/clk
      const int per_mma = (t.bn >> 1) > 40 ? (t.bn >> 1) : 40;
      const int mma = kb_unit * 4 * terms * per_mma;
      const int cost = rounds * ((ingest > mma ? ingest : mma) + 1500) + (ks > 1 ? 300 * ks : 0);  // + the reduction's reads
      if (cost < best_cost) {
        best_cost = cost;
        best.idx = ci;
        best.tiles_m = tm;
        best.tiles_n = tn;
        best.ksplit = ks;
      }
    }
  }
  return best;
}
__device__ __forceinline__ StackTile stack_tile_rt(int i) {  // run-time index -> table entry (selects, no memory)
  StackTile t = stack_tile(0);
#pragma unroll
  for (int c = 1; c < kStTiles; ++c)
    if (i == c) t = stack_tile(c);
  return t;
}

__device__ __forceinline__ float ld_shared_f32_raw(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void st_shared_f32_raw(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}

// pick_tile_cfg by a whole warp: lane = candidate (2 passes x 8 shapes), the cheapest wins, ties go to the earlier candidate
// -- the same choice as the sequential walk, in ~1/10 of the dependent instructions (the look-ahead warp was the last to
// reach the barrier of the short phases).
__device__ __forceinline__ TileCfg pick_tile_cfg_warp(int m, int ncols, int k_blocks, int nparts, int grid, int ksplit_max, int lane) {
  const int pass = (lane >> 3) & 1, ci = lane & 7;
  const int terms = nparts == 1 ? 1 : 3;
  const StackTile t = stack_tile_rt(ci);
  const int ks = pass == 0 ? 1 : ksplit_max;
  const bool pass_ok = pass == 0 || (ksplit_max > 1 && (k_blocks & (ksplit_max - 1)) == 0);
  const int kb_unit = pass == 0 ? k_blocks : k_blocks / (ksplit_max > 0 ? ksplit_max : 1);
  const int tm = (m + t.bm - 1) / t.bm, tn = (ncols + t.bn - 1) / t.bn;
  unsigned key = 0xffffffffu;
  if (lane < 16 && pass_ok && (kb_unit & (t.kbs - 1)) == 0) {
    const int units = tm * tn * ks;
    const int rounds = units <= grid ? 1 : (units <= 2 * grid ? 2 : (units + grid - 1) / grid);
    const int ingest = (t.bm + t.bn) * nparts * kb_unit * 2;
    const int per_mma = (t.bn >> 1) > 40 ? (t.bn >> 1) : 40;
    const int mma = kb_unit * 4 * terms * per_mma;
    const int cost = rounds * ((ingest > mma ? ingest : mma) + 1500) + (ks > 1 ? 300 * ks : 0);
    key = ((unsigned)(cost < (1 << 27) ? cost : (1 << 27) - 1) << 4) | (unsigned)lane;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned other = __shfl_xor_sync(0xffffffffu, key, o);
    key = other < key ? other : key;
  }
  const int win = (int)(key & 15u);
  TileCfg best;
  best.idx = win & 7;
  best.tiles_m = __shfl_sync(0xffffffffu, tm, win);
  best.tiles_n = __shfl_sync(0xffffffffu, tn, win);
  best.ksplit = __shfl_sync(0xffffffffu, ks, win);
  return best;
}

// ---------------------------------------------------------------- LayerNorm rows ---------
// Same arithmetic as layernorm_kernel<NV> (rowwise.cu): warp per row, row in registers, two-pass variance.
template <int NV>
__device__ __noinline__ void stack_ln_rows(const StackLnOp& o, int rows, int batch, int warp_global, int warp_stride, int lane) {
  const int dim = o.dim, nvec = dim >> 2, nparts = o.nparts;
  const int kp = round_up(dim, 64);
  for (int row = warp_global; row < rows; row += warp_stride) {
    const float4* xr = reinterpret_cast<const float4*>(o.x + (size_t)row * o.ld_x);
    float4 v[NV];
    float s = 0.f;
    {
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int idx = lane + i * 32;
        v[i] = idx < nvec ? __ldcg(xr + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
        s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
      }
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

// LayerNorm behind a split-K product (dim <= 512, no halo): x = sum of the K ranges (in order) + bias + residual is
// written back (it stays the residual stream) and normalised.  FOUR warps share a row -- a lane then has the 8 range loads
// of ONE 16-byte vector in flight instead of 32 loads queued behind each other (the warp-per-row version spent 3 us in
// four or five L2 round trips) -- and the row statistics meet in shared memory (fixed order).  Two rows per CTA at a time.
__device__ __noinline__ void stack_ln_fused_rows(const StackLnOp& o, int rows, int ksplit, uint32_t red_sm) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int group = warp >> 2, wq = warp & 3;
  const int dim = o.dim, nvec = dim >> 2, nparts = o.nparts;
  const int kp = round_up(dim, 64);
  const int idx = wq * 32 + lane;
  const bool live = idx < nvec;
  const uint32_t red = red_sm + (uint32_t)group * 32u;  // 4 partial sums + 4 partial square sums per group
  for (int row = blockIdx.x + gridDim.x * group; row < rows; row += gridDim.x * 2) {
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    if (live) {
      const float4* pr = reinterpret_cast<const float4*>(o.part + (size_t)row * o.ld_part) + idx;
      float4 t[8];
#pragma unroll
      for (int ks = 0; ks < 8; ++ks)
        t[ks] = ks < ksplit ? __ldcg(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(pr) + (size_t)ks * o.part_stride))
                            : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 bb = make_float4(0.f, 0.f, 0.f, 0.f), rr = bb;
      if (o.p_bias) bb = __ldg(reinterpret_cast<const float4*>(o.p_bias) + idx);
      if (o.p_res) rr = __ldcg(reinterpret_cast<const float4*>(o.p_res + (size_t)row * o.ld_res) + idx);
      a = t[0];
#pragma unroll
      for (int ks = 1; ks < 8; ++ks)
        if (ks < ksplit) {
          a.x += t[ks].x; a.y += t[ks].y; a.z += t[ks].z; a.w += t[ks].w;
        }
      a.x += bb.x; a.y += bb.y; a.z += bb.z; a.w += bb.w;
      a.x += rr.x; a.y += rr.y; a.z += rr.z; a.w += rr.w;
      reinterpret_cast<float4*>(const_cast<float*>(o.x) + (size_t)row * o.ld_x)[idx] = a;
    }
    const float s = warp_sum((a.x + a.y) + (a.z + a.w));
    if (lane == 0) st_shared_f32_raw(red + (uint32_t)wq * 4u, s);
    asm volatile("bar.sync %0, 128;" ::"r"(2 + group) : "memory");
    const float mean = ((ld_shared_f32_raw(red) + ld_shared_f32_raw(red + 4u)) + (ld_shared_f32_raw(red + 8u) + ld_shared_f32_raw(red + 12u))) / (float)dim;
    float q = 0.f;
    if (live) {
      const float d0 = a.x - mean, d1 = a.y - mean, d2 = a.z - mean, d3 = a.w - mean;
      q = (d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3);
    }
    q = warp_sum(q);
    if (lane == 0) st_shared_f32_raw(red + 16u + (uint32_t)wq * 4u, q);
    asm volatile("bar.sync %0, 128;" ::"r"(2 + group) : "memory");
    const float var = ((ld_shared_f32_raw(red + 16u) + ld_shared_f32_raw(red + 20u)) + (ld_shared_f32_raw(red + 24u) + ld_shared_f32_raw(red + 28u))) / (float)dim;
    const float rstd = rsqrtf(var + o.eps);
    if (live) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(o.gamma) + idx);
      const float4 b = __ldg(reinterpret_cast<const float4*>(o.beta) + idx);
      const float yy[4] = {(a.x - mean) * rstd * g.x + b.x, (a.y - mean) * rstd * g.y + b.y, (a.z - mean) * rstd * g.z + b.z,
                           (a.w - mean) * rstd * g.w + b.w};
      if (o.out_f32) reinterpret_cast<float4*>(o.out_f32 + (size_t)row * o.ld_f32)[idx] = make_float4(yy[0], yy[1], yy[2], yy[3]);
      if (o.out_split) store_split4(o.out_split + (size_t)row * (nparts * kp), kp, idx * 4, yy, nparts);
    }
    if (o.out_split) {
      for (int c = dim + wq * 32 + lane; c < kp; c += 128)
        for (int pp = 0; pp < nparts; ++pp) o.out_split[(size_t)row * (nparts * kp) + pp * kp + c] = __float2bfloat16_rn(0.f);
    }
    asm volatile("bar.sync %0, 128;" ::"r"(2 + group) : "memory");  // the statistics slots are free for the group's next row
  }
}

// LayerNorm of WIDE rows (640 < dim <= 2048: the decoder's FFN norm) with the whole CTA on one row: a warp per row walked
// 16 vectors per lane through three dependent passes (10-13 us per op at 71 rows, 1/3 of the decoder program's time).
// Row statistics meet in shared memory in a fixed order (summation order differs from layernorm_kernel<16>: last bits).
__device__ __noinline__ void stack_ln_wide_rows(const StackLnOp& o, int rows, uint32_t red_sm) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int dim = o.dim, nvec = dim >> 2, nparts = o.nparts;
  const int kp = round_up(dim, 64);
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const float4* xr = reinterpret_cast<const float4*>(o.x + (size_t)row * o.ld_x);
    float4 v[2];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int idx = tid + i * kStackThreads;
      v[i] = idx < nvec ? __ldcg(xr + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
      s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    }
    s = warp_sum(s);
    if (lane == 0) st_shared_f32_raw(red_sm + (uint32_t)warp * 4u, s);
    __syncthreads();
    float tot = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) tot += ld_shared_f32_raw(red_sm + (uint32_t)w * 4u);
    const float mean = tot / (float)dim;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int idx = tid + i * kStackThreads;
      if (idx < nvec) {
        const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
        q += (a * a + b * b) + (c * c + d * d);
      }
    }
    q = warp_sum(q);
    if (lane == 0) st_shared_f32_raw(red_sm + 32u + (uint32_t)warp * 4u, q);
    __syncthreads();
    float qt = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) qt += ld_shared_f32_raw(red_sm + 32u + (uint32_t)w * 4u);
    const float rstd = rsqrtf(qt / (float)dim + o.eps);
    __nv_bfloat16* srow = o.out_split ? o.out_split + (size_t)row * (nparts * kp) : nullptr;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int idx = tid + i * kStackThreads;
      if (idx < nvec) {
        const float4 g = __ldg(reinterpret_cast<const float4*>(o.gamma) + idx);
        const float4 b = __ldg(reinterpret_cast<const float4*>(o.beta) + idx);
        const float yy[4] = {(v[i].x - mean) * rstd * g.x + b.x, (v[i].y - mean) * rstd * g.y + b.y,
                             (v[i].z - mean) * rstd * g.z + b.z, (v[i].w - mean) * rstd * g.w + b.w};
        if (o.out_f32) reinterpret_cast<float4*>(o.out_f32 + (size_t)row * o.ld_f32)[idx] = make_float4(yy[0], yy[1], yy[2], yy[3]);
        if (srow) store_split4(srow, kp, idx * 4, yy, nparts);
      }
    }
    if (srow)
      for (int c = dim + tid; c < kp; c += kStackThreads)
        for (int pp = 0; pp < nparts; ++pp) srow[pp * kp + c] = __float2bfloat16_rn(0.f);
    __syncthreads();  // the statistics slots are free for the CTA's next row
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

__device__ __forceinline__ uint32_t ld_shared_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void st_shared_u32(uint32_t addr, uint32_t v) { asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
// four 8 x 8 b16 matrices; lane l supplies the row address of matrix l / 8, row l % 8
__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}
// D += A (16 x 16, row) * B (16 x 8, col), bf16 operands, fp32 accumulators
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], const uint32_t* b) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// ---------------------------------------------------------------- the kernel -------------
__global__ void __launch_bounds__(kStackThreads, 1)
stack_kernel(const StackOp* __restrict__ ops, const CUtensorMap* __restrict__ maps, int n_ops, const StackDims dims,
             unsigned int* __restrict__ sync_words, unsigned long long* __restrict__ trace, const int flags) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t epi_base = smem_base + kStRingBytes;
  const uint32_t bar_base = epi_base + kStEpiBytes;
  const uint32_t op_base = bar_base + kStMiscBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kStMaxStages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * kStMaxStages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * kStMaxStages + 2 + a); };
  auto att_bar = [&](int s) { return bar_base + 8u * (2 * kStMaxStages + 4 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kStMaxStages + 4 + kAtSlots);
  const uint32_t bias_sm = bar_base + 512u;  // 64 fp32
  const uint32_t item_sm = bar_base + 768u;  // attention work split: [0] query rows per item, [1] items
  const uint32_t cfg_sm = bar_base + 784u;   // 2 x {tile index, tiles_m, tiles_n, ksplit, op index}: the NEXT GEMM op's shape,
                                             // worked out by warp 7 while the current op runs (slot = GEMM ordinal & 1)
  const uint32_t att_sm = bar_base + 848u;   // next attention op: {aq, items, op index, b, q_beg, q_len, kv_beg, kv_len, rem}
  const StackOp* s_op = reinterpret_cast<const StackOp*>(smem_raw + (op_base - smem_u32(smem_raw)));

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int grid = gridDim.x;

  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < kStMaxStages; ++s) {
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

  // fb200_stack_profile: CTA 0's clock at points of every op (slot 8 * op + k: 0 start, 1 producer done, 2 first stage
  // landed, 3 last MMA committed, 4 the LAST CTA's body done, 5 stores landed, 6 phase body done, 7 barrier passed)
  const bool tracing = trace != nullptr && blockIdx.x == 0 && lane == 0;
  auto stamp = [&](int slot) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    trace[slot] = t;
  };
  // bit s: parity of the next phase to wait for on stage s's barrier -- the producer's view of `empty` (a fresh barrier
  // lets parity 1 through), the MMA thread's view of `full`; per barrier, so ops may cut the ring into different stage counts
  uint32_t par_empty = 0xffffffffu, par_full = 0u;
  int tl = 0;   // tiles so far (warps 1..5)
  uint32_t att_par = 0;  // bit s: parity of the next phase of attention slot barrier s (all threads advance it alike)
  unsigned int gen = 0;
  int prefetched = -1;  // index of the last GEMM op whose weights were sent towards L2
  int gemm_count = 0;   // GEMM ops so far (all threads)
  int last_ksplit = 1;  // K split of the last GEMM op (the LayerNorm op behind a split product sums that many ranges)
  int cfg_done = -1;    // warp 7: op index whose tile shape sits in cfg_sm
  if (tid == 0) {
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(cfg_sm + 16u), "r"(-1) : "memory");
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(cfg_sm + 48u), "r"(-1) : "memory");
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(att_sm + 8u), "r"(-1) : "memory");
  }

  if (tracing && tid == 0) trace[8 * n_ops] = (unsigned long long)clock64();  // SM clock over the run (with the time stamps: MHz)
  constexpr int kOpVecs = (int)(sizeof(StackOp) / 16);
  uint4 op_next = make_uint4(0u, 0u, 0u, 0u);  // the next op record, fetched while the current op runs
  if (tid < kOpVecs) op_next = __ldg(reinterpret_cast<const uint4*>(ops) + tid);
  for (int oi = 0; oi < n_ops; ++oi) {
    if (tid < kOpVecs) reinterpret_cast<uint4*>(smem_raw + (op_base - smem_u32(smem_raw)))[tid] = op_next;
    __syncthreads();
    if (tid < kOpVecs && oi + 1 < n_ops) op_next = __ldg(reinterpret_cast<const uint4*>(ops + oi + 1) + tid);
    if (tracing && tid == 0) stamp(8 * oi);
    const StackOp& op = *s_op;
    const int rows = op.rows_slot >= 0 ? dims.v[op.rows_slot] : op.rows;
    const int batch = op.batch_slot >= 0 ? dims.v[op.batch_slot] : op.batch;

    // towards L2 while this op runs (hints; nothing waits for them): the weights of the next GEMM op, and the small
    // parameter vectors (bias, LayerNorm gamma / beta, FSMN taps) of the next op -- every one of them is a cold HBM line
    // on a path that streams 880 MB of weights per request
    if (warp == 7 && !(flags & 1)) {
      auto l2_prefetch = [&](const void* base, unsigned long long total) {
        const char* b = (const char*)base;
        const unsigned long long chunk = 4096;
        for (unsigned long long off = ((unsigned long long)blockIdx.x * 32 + lane) * chunk; off < total;
             off += (unsigned long long)grid * 32 * chunk) {
          const unsigned long long left = total - off;
          const uint32_t nbytes = (uint32_t)(left < chunk ? (left & ~15ull) : chunk);
          if (nbytes) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(b + off), "r"(nbytes) : "memory");
        }
      };
      if (op.next_gemm > prefetched && op.next_gemm < n_ops) {
        const StackOp* nx = ops + op.next_gemm;
        l2_prefetch(nx->u.g.w_base, nx->u.g.w_bytes);
        if (nx->u.g.bias) l2_prefetch(nx->u.g.bias, (unsigned long long)nx->u.g.n * 4ull);
        if (lane < kStMapsPerGemm) asm volatile("prefetch.tensormap [%0];" ::"l"(maps + nx->u.g.map0 + lane) : "memory");
      }
      if (oi + 1 < n_ops) {
        const StackOp* nx = ops + oi + 1;
        const int kind = nx->kind;
        if (kind == kStLayerNorm) {
          l2_prefetch(nx->u.ln.gamma, (unsigned long long)nx->u.ln.dim * 4ull);
          l2_prefetch(nx->u.ln.beta, (unsigned long long)nx->u.ln.dim * 4ull);
          if (nx->u.ln.p_bias) l2_prefetch(nx->u.ln.p_bias, (unsigned long long)nx->u.ln.dim * 4ull);
        } else if (kind == kStFsmn) {
          l2_prefetch(nx->u.f.weight, (unsigned long long)nx->u.f.dim * nx->u.f.ksize * 4ull);
        } else if (kind == kStAttention && lane < 2) {
          asm volatile("prefetch.tensormap [%0];" ::"l"(maps + nx->u.a.map0 + lane) : "memory");
        }
      }
    }
    if (op.next_gemm > prefetched && op.next_gemm < n_ops) prefetched = op.next_gemm;
    // look-ahead on an otherwise idle thread: the tile shape of the next GEMM op (a few hundred dependent instructions
    // that used to sit in front of every product's first load) and the work split of an attention op right behind this one
    if (warp == 7) {
      const int target = op.next_gemm;
      if (target < n_ops && target != cfg_done) {  // (warp-uniform)
        const StackOp* nx = ops + target;
        const int nrows = nx->rows_slot >= 0 ? dims.v[nx->rows_slot] : nx->rows;
        const int ncols_n = nx->u.g.out_split ? nx->u.g.out_kp : nx->u.g.n;
        const TileCfg c = pick_tile_cfg_warp(nrows, ncols_n, nx->u.g.taps * nx->u.g.k_blocks, nx->u.g.nparts, grid,
                                             nx->u.g.taps == 1 ? nx->u.g.ksplit_max : 1, lane);
        if (lane == 0) {
          const uint32_t slot = cfg_sm + (uint32_t)((gemm_count + (op.kind == kStGemm ? 1 : 0)) & 1) * 32u;
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(slot), "r"(c.idx), "r"(c.tiles_m), "r"(c.tiles_n), "r"(c.ksplit) : "memory");
          asm volatile("st.shared.b32 [%0], %1;" ::"r"(slot + 16u), "r"(target) : "memory");
        }
        cfg_done = target;
      }
    }
    if (warp == 7 && lane == 0) {
      if (oi + 1 < n_ops && ops[oi + 1].kind == kStAttention) {
        const StackOp* nx = ops + oi + 1;
        const StackAttnOp& na = nx->u.a;
        const int nb = nx->batch_slot >= 0 ? dims.v[nx->batch_slot] : nx->batch;
        const int total_q = __ldg(na.cu_q + nb);
        int aq = (na.heads * total_q + grid - 1) / grid;
        aq = aq < 1 ? 1 : (aq > kAtRows ? kAtRows : aq);
        int items = 0;
        for (;;) {
          items = 0;
          for (int b = 0; b < nb; ++b) {
            const int len = __ldg(na.cu_q + b + 1) - __ldg(na.cu_q + b);
            const int kv = __ldg(na.cu_kv + b + 1) - __ldg(na.cu_kv + b);
            if (kv > 0) items += ((len + aq - 1) / aq) * na.heads;
          }
          if (items <= grid || aq >= kAtRows) break;
          ++aq;
        }
        int b = 0, rem = blockIdx.x, q_beg = 0, q_len = 0, kv_beg = 0, kv_len = 0;
        for (; b < nb; ++b) {
          q_beg = __ldg(na.cu_q + b);
          q_len = __ldg(na.cu_q + b + 1) - q_beg;
          kv_beg = __ldg(na.cu_kv + b);
          kv_len = __ldg(na.cu_kv + b + 1) - kv_beg;
          const int n_it = kv_len > 0 ? ((q_len + aq - 1) / aq) * na.heads : 0;
          if (rem < n_it) break;
          rem -= n_it;
        }
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(att_sm), "r"(aq), "r"(items), "r"(oi + 1), "r"(b) : "memory");
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(att_sm + 16u), "r"(q_beg), "r"(q_len), "r"(kv_beg), "r"(kv_len) : "memory");
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(att_sm + 32u), "r"(rem) : "memory");
      }
    }

    if (op.kind == kStGemm) {
      // ================================================================ GEMM ============
      const StackGemmOp& g = op.u.g;
      const int m = rows, n = g.n, nparts = g.nparts;
      const int ncols = g.out_split ? g.out_kp : n;
      TileCfg cfg;
      {
        const uint32_t slot = cfg_sm + (uint32_t)(gemm_count & 1) * 32u;
        int tag;
        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(cfg.idx), "=r"(cfg.tiles_m), "=r"(cfg.tiles_n), "=r"(cfg.ksplit) : "r"(slot) : "memory");
        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tag) : "r"(slot + 16u) : "memory");
        if (tag != oi) cfg = pick_tile_cfg(m, ncols, g.taps * g.k_blocks, nparts, grid, g.taps == 1 ? g.ksplit_max : 1);
      }
      ++gemm_count;
      last_ksplit = cfg.ksplit;
      const int ksplit = cfg.ksplit, kb_unit = g.k_blocks / ksplit;  // unit = (tile, K range)
      const StackTile tc = stack_tile_rt(cfg.idx);
      const int bm = tc.bm, bn = tc.bn, kbs = tc.kbs;
      const int total_tiles = cfg.tiles_m * cfg.tiles_n * ksplit;  // units
      const CUtensorMap* map_f32 = maps + g.map0;
      const CUtensorMap* map_split = maps + g.map0 + 1;
      const CUtensorMap* map_a = maps + g.map0 + kStAMap0 + tc.a_map;
      const CUtensorMap* map_w = maps + g.map0 + kStWMap0 + tc.w_map;
      const uint32_t a_blk = (uint32_t)bm * 128u, b_blk = (uint32_t)bn * 128u;  // one [rows x 64] block of one part
      const uint32_t a_bytes = (uint32_t)(nparts * kbs) * a_blk, b_bytes = (uint32_t)(nparts * kbs) * b_blk;
      const uint32_t stage_bytes = a_bytes + b_bytes;
      const int n_stages = (int)(kStRingBytes / stage_bytes) < kStMaxStages ? (int)(kStRingBytes / stage_bytes) : kStMaxStages;

      if (warp == 0 || warp == 6) {
        // ===== TMA producers: warp 0 brings the A blocks (and arms the barrier for the whole stage), warp 6 the W blocks --
        // two issuing warps because an instruction occupies its warp for ~200 cycles
        // (elect.sync, not `lane == 0`: the compiler then keeps descriptors and barrier addresses in uniform registers --
        // with a lane test every UTMALDG / UTCHMMA was wrapped in an ELECT + 7 x R2UR loop, ~100 cycles apiece)
        const bool leader = elect_one();
        if (leader) {
          const bool a_role = warp == 0;
          int s = 0;
          for (int unit = blockIdx.x; unit < total_tiles; unit += grid) {
            const int tile = unit / ksplit, kb_lo = (unit - tile * ksplit) * kb_unit;
            const int m0 = (tile / cfg.tiles_n) * bm, n0 = (tile % cfg.tiles_n) * bn;
            for (int tap = 0; tap < g.taps; ++tap) {
              for (int kb0 = kb_lo; kb0 < kb_lo + kb_unit; kb0 += kbs) {
                mbar_wait(empty_bar(s), (par_empty >> s) & 1u);
                par_empty ^= 1u << s;
                const uint32_t sa = smem_base + (uint32_t)s * stage_bytes;
                if (a_role) {
                  mbar_expect_tx(full_bar(s), stage_bytes);
                  tma_load_4d(sa, map_a, 0, m0 + g.tap_row[tap], kb0, 0, full_bar(s));
                } else {
                  tma_load_4d(sa + a_bytes, map_w, 0, tap * n + n0, kb0, 0, full_bar(s));
                }
                s = s + 1 == n_stages ? 0 : s + 1;
              }
            }
          }
        }
        if (leader && trace != nullptr && blockIdx.x == 0 && warp == 0) stamp(8 * oi + 1);
        {  // the barrier parities live in every lane: take the elected lane's copy
          const unsigned who = __ballot_sync(0xffffffffu, leader);
          par_empty = __shfl_sync(0xffffffffu, par_empty, __ffs(who) - 1);
        }
      } else if (warp == 1) {
        const bool leader = elect_one();
        if (leader) {  // ===== MMA issuer =====
          const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
          int s = 0, t = tl;
          bool first = true;
          for (int unit = blockIdx.x; unit < total_tiles; unit += grid, ++t) {
            const int kb_lo = (unit % ksplit) * kb_unit;
            const int acc = t & 1;
            mbar_wait(tempty_bar(acc), (((uint32_t)t >> 1) & 1u) ^ 1u);
            tcgen05_fence_after();
            const uint32_t tmem_d = tmem_base + (uint32_t)(acc * 64);
            bool fresh = true;  // the tile's first MMA overwrites the accumulator
            for (int tap = 0; tap < g.taps; ++tap) {
              for (int kb0 = kb_lo; kb0 < kb_lo + kb_unit; kb0 += kbs) {
                mbar_wait(full_bar(s), (par_full >> s) & 1u);
                par_full ^= 1u << s;
                if (trace != nullptr && blockIdx.x == 0 && first) stamp(8 * oi + 2);
                first = false;
                tcgen05_fence_after();
                const uint32_t sa = smem_base + (uint32_t)s * stage_bytes;
                const uint32_t sb = sa + a_bytes;
                const int kb_here = kb_lo + kb_unit - kb0 < kbs ? kb_lo + kb_unit - kb0 : kbs;
                // The issue loop is the critical path of a small-tile product (one thread, ~100 dependent-instruction cycles
                // per MMA when each descriptor is rebuilt): descriptors of a k-block are built once (hi / mid of A and W)
                // and stepped by adding the 16-byte-unit offset of the next 16-element K slice to their address field.
                for (int kb = 0; kb < ((flags & 2) ? 0 : kb_here); ++kb) {  // flags bit 1: timing experiment without the MMAs
                  const uint64_t da0 = umma_desc_sw128(sa + (uint32_t)kb * a_blk), db0 = umma_desc_sw128(sb + (uint32_t)kb * b_blk);
                  if (nparts == 2 && !(flags & 16)) {  // flags bit 4: timing experiment with one term per k-block
                    const uint64_t da1 = da0 + (uint64_t)((uint32_t)kbs * a_blk >> 4), db1 = db0 + (uint64_t)((uint32_t)kbs * b_blk >> 4);
#pragma unroll
                    for (int k = 0; k < SBK / 16; ++k) umma_bf16(tmem_d, da0 + 2 * k, db1 + 2 * k, idesc, (fresh && k == 0) ? 0u : 1u);
#pragma unroll
                    for (int k = 0; k < SBK / 16; ++k) umma_bf16(tmem_d, da1 + 2 * k, db0 + 2 * k, idesc, 1u);
#pragma unroll
                    for (int k = 0; k < SBK / 16; ++k) umma_bf16(tmem_d, da0 + 2 * k, db0 + 2 * k, idesc, 1u);
                  } else {
#pragma unroll
                    for (int k = 0; k < SBK / 16; ++k) umma_bf16(tmem_d, da0 + 2 * k, db0 + 2 * k, idesc, (fresh && k == 0) ? 0u : 1u);
                  }
                  fresh = false;
                }
                umma_commit(empty_bar(s));
                s = s + 1 == n_stages ? 0 : s + 1;
              }
            }
            umma_commit(tfull_bar(acc));
          }
          if (trace != nullptr && blockIdx.x == 0) stamp(8 * oi + 3);
        }
        for (int unit = blockIdx.x; unit < total_tiles; unit += grid) ++tl;  // (every lane keeps the count)
        {
          const unsigned who = __ballot_sync(0xffffffffu, leader);
          par_full = __shfl_sync(0xffffffffu, par_full, __ffs(who) - 1);
        }
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
        for (int unit = blockIdx.x; unit < total_tiles; unit += grid, ++t) {
          const int tile = unit / ksplit;
          const int part_r0 = (unit - tile * ksplit) * g.part_rows;  // split-K: this K range's rows of the scratch matrix
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
            const bool do_f32 = has_f32 && col0 >= f32_col0;
            if (!(flags & 32)) {
              // plain stores: a thread owns 32 consecutive columns of its row (128 B of fp32, 64 B per bf16 part).  The
              // TMA-store version (staging boxes + cp.async.bulk) published its bytes ~1.2 us later at the barrier.
              if (row_ok) {
                if (do_f32) {
                  float* dst = g.out_f32 + (size_t)(part_r0 + row) * g.ld_out + (col0 - f32_col0);
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
                  __nv_bfloat16* dst = g.out_split + (size_t)row * (out_parts * out_kp) + col0;
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
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncwarp();
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
            if (elect_one()) {
              const int r0 = m0 + lg * 32;
              if (do_f32) tma_store_2d(map_f32, fbuf, col0 - f32_col0, part_r0 + r0);
              if (has_split) {
                tma_store_2d(map_split, sbuf, col0, r0);
                if (out_parts > 1) tma_store_2d(map_split, sbuf + 2048u, out_kp + col0, r0);
              }
              asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
          }
        }
        tl = t;
        if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // this warp's stores have landed
        if (tracing && warp == 2) stamp(8 * oi + 5);
        if (trace != nullptr && warp == 2 && lane == 0) {  // slot 4 of a GEMM op: the LAST CTA's stores landed
          unsigned long long tt;
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tt));
          atomicMax(trace + 8 * oi + 4, tt);
        }
        __syncwarp();
      }
    } else if (op.kind == kStLayerNorm) {
      // ================================================================ LayerNorm =======
      const StackLnOp& o = op.u.ln;
      const int wg = blockIdx.x + grid * warp, ws = grid * (kStackThreads / 32);
      const int ksplit = o.part != nullptr ? last_ksplit : 1;  // the split-K GEMM op right before this op chose it
      // (the recorder fuses a split-K reduction only into LayerNorms of at most 512 columns without the halo layout)
      if (o.part != nullptr) stack_ln_fused_rows(o, rows, ksplit, bias_sm);
      else if (o.dim <= 512) stack_ln_rows<4>(o, rows, batch, wg, ws, lane);
      else if (o.dim <= 640) stack_ln_rows<5>(o, rows, batch, wg, ws, lane);
      else if (!o.halo && rows <= 2 * grid) stack_ln_wide_rows(o, rows, bias_sm);  // whole CTA per row
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
      int aq, items, pre_tag, pre_b, pre_qb, pre_ql, pre_kb, pre_kl, pre_rem;
      asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(aq), "=r"(items), "=r"(pre_tag), "=r"(pre_b) : "r"(att_sm) : "memory");
      asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(pre_qb), "=r"(pre_ql), "=r"(pre_kb), "=r"(pre_kl) : "r"(att_sm + 16u) : "memory");
      asm volatile("ld.shared.b32 %0, [%1];" : "=r"(pre_rem) : "r"(att_sm + 32u) : "memory");
      if (pre_tag != oi) {  // not looked ahead (first op of a program): work it out here
        if (tid == 0) {  // query rows per item: the smallest AQ whose item count fits one round (capped at 8)
          const int total_q = __ldg(o.cu_q + batch);
          int a = (heads * total_q + grid - 1) / grid;
          a = a < 1 ? 1 : (a > kAtRows ? kAtRows : a);
          int n_items = 0;
          for (;;) {
            n_items = 0;
            for (int b = 0; b < batch; ++b) {
              const int len = __ldg(o.cu_q + b + 1) - __ldg(o.cu_q + b);
              const int kv = __ldg(o.cu_kv + b + 1) - __ldg(o.cu_kv + b);
              if (kv > 0) n_items += ((len + a - 1) / a) * heads;
            }
            if (n_items <= grid || a >= kAtRows) break;
            ++a;
          }
          asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(item_sm), "r"(a), "r"(n_items) : "memory");
        }
        __syncthreads();
        asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(aq), "=r"(items) : "r"(item_sm) : "memory");
      }
      for (int item = blockIdx.x; item < items; item += grid) {
        // item -> (utterance, chunk, head)
        int b = 0, rem = item, q_beg = 0, q_len = 0, kv_beg = 0, kv_len = 0;
        if (pre_tag == oi && item == (int)blockIdx.x) {  // this CTA's first item was decoded ahead of time
          b = pre_b; rem = pre_rem; q_beg = pre_qb; q_len = pre_ql; kv_beg = pre_kb; kv_len = pre_kl;
        } else
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
        const CUtensorMap* map_k = maps + o.map0;
        const CUtensorMap* map_v = maps + o.map0 + 1;
        auto issue_tile = [&](int t) {  // warps 0 / 1: the K / V blocks of tile t into slot t % 3, ONE instruction each
          if (elect_one()) {              // (box = [part][half of the head dim][64 keys][64 dims])
            const int s = t % kAtSlots;
            const uint32_t slot = smem_base + (uint32_t)s * kAtSlot;
            const int row = kv_beg + t * kAtTile;
            if (warp == 0) {
              mbar_expect_tx(att_bar(s), (uint32_t)(nparts * 4) * kAtBox);
              tma_load_4d(slot, map_k, 0, row, head * 2, 0, att_bar(s));
            } else {
              tma_load_4d(slot + 2 * kAtPart, map_v, 0, row, head * 2, 0, att_bar(s));
            }
          }
          __syncwarp();
        };
        if (warp < 2)
          for (int t = 0; t < ntiles && t < kAtSlots; ++t) issue_tile(t);
        // ---- query rows as they lie in the split matrix (bf16 hi | mid): 16-byte chunks into the swizzled staging tile
        for (int id = tid; id < 2 * kAtRows * 16; id += kStackThreads) {
          const int pp = id >> 8, r = (id >> 4) & 15, c = id & 15;
          uint4 val = make_uint4(0u, 0u, 0u, 0u);
          if (r < nq && pp < nparts)
            val = __ldcg(reinterpret_cast<const uint4*>(o.q + (size_t)(q_beg + q0 + r) * o.ld_q + hoff + (size_t)pp * o.q_ps + c * 8));
          st_shared_v4(smem_base + kAtQ + (uint32_t)pp * 4096u + (uint32_t)r * 256u + (uint32_t)((c ^ (r & 7)) << 4), val.x, val.y, val.z, val.w);
        }
        __syncthreads();

        if (tracing && tid == 0 && item == (int)blockIdx.x) stamp(8 * oi + 1);
        // m16n8k16 fragments: g = row / key-in-tile index, tq = column pair
        const int g = lane >> 2, tq = lane & 3;
        uint32_t qa[2][8][4];  // A fragments of Q: [part][k-step of 16 dims]
        {
          const int r = (lane & 7) + ((lane >> 3) & 1) * 8;
#pragma unroll
          for (int pp = 0; pp < 2; ++pp)
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
              const int c = ks * 2 + (lane >> 4);
              ldmatrix_x4(smem_base + kAtQ + (uint32_t)pp * 4096u + (uint32_t)r * 256u + (uint32_t)((c ^ (r & 7)) << 4), qa[pp][ks]);
            }
        }
        float oacc[2][4];  // O: rows g / g + 8, dims warp * 16 + j * 8 + 2 tq (+1)
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
          for (int u = 0; u < 4; ++u) oacc[j][u] = 0.f;
        float m_run[2] = {-INFINITY, -INFINITY};  // rows g, g + 8 (every thread of a row holds the same values)
        float l_part[2] = {0.f, 0.f};             // this thread's share of the row sums (its two keys of every tile)

        for (int t = 0; t < ntiles; ++t) {
          const int s = t % kAtSlots;
          const int nvalid = kv_len - t * kAtTile < kAtTile ? kv_len - t * kAtTile : kAtTile;
          mbar_wait(att_bar(s), (att_par >> s) & 1u);
          att_par ^= 1u << s;
          if (tracing && tid == 0 && item == (int)blockIdx.x && t == 0) stamp(8 * oi + 2);
          const uint32_t slot = smem_base + (uint32_t)s * kAtSlot;
          // ---- S[16 x 8 keys of this warp] = Q K^T over 128 dims: hi.mid + mid.hi + hi.hi per 16-dim step
          // (one accumulator per cross term: three independent MMA chains of 8 instead of one of 24 -- the phase is bound
          // by the MMA latency, not by its rate)
          float sc[4] = {0.f, 0.f, 0.f, 0.f}, sc1[4] = {0.f, 0.f, 0.f, 0.f}, sc2[4] = {0.f, 0.f, 0.f, 0.f};
          {
            const int key = warp * 8 + g;
            const uint32_t krow = slot + (uint32_t)key * 128u + (uint32_t)tq * 4u;
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
              const uint32_t kb = krow + (uint32_t)(ks >> 2) * kAtBox;
              const uint32_t c0 = (uint32_t)((((ks & 3) * 2) ^ (key & 7)) << 4), c1 = (uint32_t)((((ks & 3) * 2 + 1) ^ (key & 7)) << 4);
              uint32_t bh[2], bm[2] = {0u, 0u};
              bh[0] = ld_shared_u32(kb + c0);
              bh[1] = ld_shared_u32(kb + c1);
              if (nparts > 1) {
                bm[0] = ld_shared_u32(kb + kAtPart + c0);
                bm[1] = ld_shared_u32(kb + kAtPart + c1);
                mma_bf16_16816(sc1, qa[0][ks], bm);
                mma_bf16_16816(sc2, qa[1][ks], bh);
              }
              mma_bf16_16816(sc, qa[0][ks], bh);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) sc[u] = (sc1[u] + sc2[u]) + sc[u];  // small terms first
          }
          // ---- row maxima across the tile: quad, then the eight warps through shared memory
          const int k0 = warp * 8 + tq * 2;
          const bool v0 = k0 < nvalid, v1 = k0 + 1 < nvalid;
          sc[0] = v0 ? sc[0] * o.scale : -INFINITY;
          sc[1] = v1 ? sc[1] * o.scale : -INFINITY;
          sc[2] = v0 ? sc[2] * o.scale : -INFINITY;
          sc[3] = v1 ? sc[3] * o.scale : -INFINITY;
          {
            float mx0 = fmaxf(sc[0], sc[1]), mx1 = fmaxf(sc[2], sc[3]);
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
            if (tq == 0) {
              st_shared_f32(smem_base + kAtStat + (uint32_t)(warp * 16 + g) * 4u, mx0);
              st_shared_f32(smem_base + kAtStat + (uint32_t)(warp * 16 + g + 8) * 4u, mx1);
            }
          }
          // (the last tile's value rows past the utterance may hold anything: zero them so that 0 * garbage stays 0)
          if (nvalid < kAtTile) {
            for (int id = tid; id < (kAtTile - nvalid) * 16 * nparts; id += kStackThreads) {
              const int pp = id / ((kAtTile - nvalid) * 16), rem2 = id - pp * (kAtTile - nvalid) * 16;
              const int r = nvalid + (rem2 >> 4), c = rem2 & 15;
              st_shared_v4(slot + 2 * kAtPart + (uint32_t)pp * kAtPart + (uint32_t)(c >> 3) * kAtBox + (uint32_t)r * 128u + (uint32_t)((c & 7) << 4), 0u, 0u, 0u, 0u);
            }
          }
          __syncthreads();
          float corr[2];
          {
            float mn0 = m_run[0], mn1 = m_run[1];
#pragma unroll
            for (int w = 0; w < 8; ++w) {
              mn0 = fmaxf(mn0, ld_shared_f32(smem_base + kAtStat + (uint32_t)(w * 16 + g) * 4u));
              mn1 = fmaxf(mn1, ld_shared_f32(smem_base + kAtStat + (uint32_t)(w * 16 + g + 8) * 4u));
            }
            corr[0] = expf(m_run[0] - mn0);  // exp(-inf) = 0 on the first tile
            corr[1] = expf(m_run[1] - mn1);
            m_run[0] = mn0;
            m_run[1] = mn1;
          }
          // ---- probabilities of this warp's keys -> bf16 hi / mid in the staging tile (the A operand of P V)
          {
            const float p0 = v0 ? expf(sc[0] - m_run[0]) : 0.f, p1 = v1 ? expf(sc[1] - m_run[0]) : 0.f;
            const float p2 = v0 ? expf(sc[2] - m_run[1]) : 0.f, p3 = v1 ? expf(sc[3] - m_run[1]) : 0.f;
            l_part[0] = l_part[0] * corr[0] + (p0 + p1);
            l_part[1] = l_part[1] * corr[1] + (p2 + p3);
            const __nv_bfloat162 h01 = __floats2bfloat162_rn(p0, p1), h23 = __floats2bfloat162_rn(p2, p3);
            const __nv_bfloat162 m01 = __floats2bfloat162_rn(p0 - __low2float(h01), p1 - __high2float(h01));
            const __nv_bfloat162 m23 = __floats2bfloat162_rn(p2 - __low2float(h23), p3 - __high2float(h23));
            const uint32_t pa0 = smem_base + kAtP + (uint32_t)g * 128u + (uint32_t)((warp ^ (g & 7)) << 4) + (uint32_t)tq * 4u;
            const uint32_t pa1 = smem_base + kAtP + (uint32_t)(g + 8) * 128u + (uint32_t)((warp ^ (g & 7)) << 4) + (uint32_t)tq * 4u;
            st_shared_u32(pa0, *reinterpret_cast<const uint32_t*>(&h01));
            st_shared_u32(pa1, *reinterpret_cast<const uint32_t*>(&h23));
            st_shared_u32(pa0 + 2048u, *reinterpret_cast<const uint32_t*>(&m01));
            st_shared_u32(pa1 + 2048u, *reinterpret_cast<const uint32_t*>(&m23));
          }
          __syncthreads();
          // ---- O[16 x 16 dims of this warp] = O * corr + P V over the tile's 64 keys
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            oacc[j][0] *= corr[0]; oacc[j][1] *= corr[0];
            oacc[j][2] *= corr[1]; oacc[j][3] *= corr[1];
          }
          {
            const int pr = (lane & 7) + ((lane >> 3) & 1) * 8;          // P: ldmatrix row of this lane
            const int mi = lane >> 3;                                   // V: matrix index of this lane
            const int vkey = ((mi & 1) << 3) + (lane & 7);              //    key within the 16-key step
            const int vdim = warp * 16 + ((mi >> 1) << 3);              //    first dim of its 8 x 8 block
            const uint32_t vbase = slot + 2 * kAtPart + (uint32_t)(vdim >> 6) * kAtBox;
            const int vchunk = (vdim & 63) >> 3;
            float t1[2][4], t2[2][4];  // the hi.mid and mid.hi terms of this tile (independent MMA chains)
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
              for (int u = 0; u < 4; ++u) t1[j][u] = t2[j][u] = 0.f;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              uint32_t ph[4], pm[4] = {0u, 0u, 0u, 0u}, vh[4], vm[4] = {0u, 0u, 0u, 0u};
              const int pc = kk * 2 + (lane >> 4);
              const uint32_t paddr = smem_base + kAtP + (uint32_t)pr * 128u + (uint32_t)((pc ^ (pr & 7)) << 4);
              const int key = kk * 16 + vkey;
              const uint32_t vaddr = vbase + (uint32_t)key * 128u + (uint32_t)((vchunk ^ (key & 7)) << 4);
              ldmatrix_x4(paddr, ph);
              ldmatrix_x4_trans(vaddr, vh);
              if (nparts > 1) {
                ldmatrix_x4(paddr + 2048u, pm);
                ldmatrix_x4_trans(vaddr + kAtPart, vm);
                mma_bf16_16816(t1[0], ph, vm);
                mma_bf16_16816(t1[1], ph, vm + 2);
                mma_bf16_16816(t2[0], pm, vh);
                mma_bf16_16816(t2[1], pm, vh + 2);
              }
              mma_bf16_16816(oacc[0], ph, vh);
              mma_bf16_16816(oacc[1], ph, vh + 2);
            }
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
              for (int u = 0; u < 4; ++u) oacc[j][u] += t1[j][u] + t2[j][u];
          }
          __syncthreads();  // tile consumed: its slot, the row maxima and P may be overwritten
          if (warp < 2 && t + kAtSlots < ntiles) issue_tile(t + kAtSlots);
        }
        if (tracing && tid == 0 && item == (int)blockIdx.x) stamp(8 * oi + 3);
        // ---- row sums: quad, then the eight warps (fixed order); normalise, store
        {
          float l0 = l_part[0], l1 = l_part[1];
          l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
          l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
          l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
          l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
          if (tq == 0) {
            st_shared_f32(smem_base + kAtStat + (uint32_t)(warp * 16 + g) * 4u, l0);
            st_shared_f32(smem_base + kAtStat + (uint32_t)(warp * 16 + g + 8) * 4u, l1);
          }
        }
        __syncthreads();
        {
          float l0 = 0.f, l1 = 0.f;
#pragma unroll
          for (int w = 0; w < 8; ++w) {
            l0 += ld_shared_f32(smem_base + kAtStat + (uint32_t)(w * 16 + g) * 4u);
            l1 += ld_shared_f32(smem_base + kAtStat + (uint32_t)(w * 16 + g + 8) * 4u);
          }
          const float inv0 = 1.f / l0, inv1 = 1.f / l1;
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const int col = hoff + warp * 16 + j * 8 + tq * 2;
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
              const int r = g + hf * 8;
              if (r < nq) {
                const float y0 = oacc[j][hf * 2] * (hf ? inv1 : inv0), y1 = oacc[j][hf * 2 + 1] * (hf ? inv1 : inv0);
                const size_t row = (size_t)(q_beg + q0 + r);
                if (o.out_f32) *reinterpret_cast<float2*>(o.out_f32 + row * o.ld_out + col) = make_float2(y0, y1);
                if (o.out_split) {
                  __nv_bfloat16* srow = o.out_split + row * (size_t)(o.out_parts * o.out_kp);
                  const __nv_bfloat162 h = __floats2bfloat162_rn(y0, y1);
                  *reinterpret_cast<__nv_bfloat162*>(srow + col) = h;
                  if (o.out_parts > 1)
                    *reinterpret_cast<__nv_bfloat162*>(srow + o.out_kp + col) = __floats2bfloat162_rn(y0 - __low2float(h), y1 - __high2float(h));
                }
              }
            }
          }
        }
        __syncthreads();  // the staging tiles and the statistics are free for the next item
      }
    }

    // ---- end of the op
    if (tracing && tid == 0) stamp(8 * oi + 6);
    if (trace != nullptr && tid == 0 && op.kind != kStGemm) {  // slot 4: the LAST CTA's arrival at the end of the op body
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      atomicMax(trace + 8 * oi + 4, t);
    }
    if (op.barrier_after)
      grid_barrier(sync_words, gen, flags);
    else
      __syncthreads();
    if (tracing && tid == 0) stamp(8 * oi + 7);
  }

  if (tracing && tid == 0) trace[8 * n_ops + 1] = (unsigned long long)clock64();
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
  unsigned long long* d_trace = nullptr;  // fb200_stack_profile(handle, 1): n_ops + 1 time stamps of the last run
  float* d_scratch[2] = {nullptr, nullptr};  // split-K partial sums (two buffers, taken in turn)
  int fused = 0;
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
  g.ksplit_max = 1;
  g.part_rows = 0;
  g.bias = p.bias;
  g.res1 = p.res1;
  g.out_f32 = p.out_f32;
  g.out_split = p.out_split;
  g.w_base = (const char*)w;
  g.w_bytes = (unsigned long long)w_rows * ldw * 2ull;
  // output boxes as in gemm_host.inc; operand maps for every tile shape the kernel may pick at run time
  CUtensorMap mf, ms, mo;
  int rc = make_tma_map(&mo, a, 2, (uint64_t)a_rows, (uint64_t)lda, (uint64_t)lda, 64, 32, 128);  // placeholder for absent outputs
  if (rc) return rc;
  mf = mo;
  ms = mo;
  if (p.out_f32) {
    rc = make_tma_map(&mf, p.out_f32, 4, (uint64_t)p.m, (uint64_t)(p.n - p.f32_col0), (uint64_t)p.ld_out, 32, 32, 128);
    if (rc) return rc;
  }
  if (p.out_split) {
    rc = make_tma_map(&ms, p.out_split, 2, (uint64_t)p.m, (uint64_t)p.out_parts * p.out_kp, (uint64_t)p.out_parts * p.out_kp, 32,
                      32, 64);
    if (rc) return rc;
  }
  pr.maps.push_back(mf);
  pr.maps.push_back(ms);
  for (int i = 0; i < 6; ++i) {
    CUtensorMap ma;
    rc = make_tma_map_operand4(&ma, (const __nv_bfloat16*)a + p.a_col0, (uint64_t)a_rows, (uint64_t)lda, (uint64_t)p.k_blocks,
                               (uint64_t)a_parts, (uint64_t)p.kp_a, (uint32_t)kHostAMaps[i][0], (uint32_t)kHostAMaps[i][1]);
    if (rc) return rc;
    pr.maps.push_back(ma);
  }
  for (int i = 0; i < 5; ++i) {
    CUtensorMap mw;
    rc = make_tma_map_operand4(&mw, w, (uint64_t)w_rows, (uint64_t)ldw, (uint64_t)p.k_blocks, (uint64_t)a_parts, (uint64_t)p.kp_w,
                               (uint32_t)kHostWMaps[i][0], (uint32_t)kHostWMaps[i][1]);
    if (rc) return rc;
    pr.maps.push_back(mw);
  }
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
  o.k = (const __nv_bfloat16*)k;
  o.v = (const __nv_bfloat16*)v;
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
  o.map0 = (int)pr.maps.size();
  o.k_col0 = k_col0;
  o.v_col0 = v_col0;
  CUtensorMap mk, mv;  // box = both 64-dim halves of one head, every part, 64 keys
  int rc = make_tma_map_operand4(&mk, (const __nv_bfloat16*)k + k_col0, (uint64_t)kv_rows, (uint64_t)ld_k, (uint64_t)heads * 2,
                                 (uint64_t)nparts, (uint64_t)kv_part_stride, 64, 2);
  if (rc) return rc;
  rc = make_tma_map_operand4(&mv, (const __nv_bfloat16*)v + v_col0, (uint64_t)kv_rows, (uint64_t)ld_v, (uint64_t)heads * 2,
                             (uint64_t)nparts, (uint64_t)kv_part_stride, 64, 2);
  if (rc) return rc;
  pr.maps.push_back(mk);
  pr.maps.push_back(mv);
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
  // Split-K fusion: a product whose only consumer is the LayerNorm op right behind it (out-projection -> norm2, FFN w2 ->
  // the next layer's norm1, ...) may be cut into K ranges -- what bounds a small-tile product is the NUMBER of MMAs its
  // one issuing thread walks through -- with the LayerNorm op summing the ranges (+ bias, + residual) as it reads its
  // rows.  Both ops are separated from their neighbours by barriers, so the scratch matrix is never read and written in
  // the same phase; two scratch buffers are used in turn all the same.
  static int split_on = -1;
  if (split_on < 0) {
    const char* e = getenv("FB200_STACK_SPLITK");
    split_on = e ? atoi(e) : 1;
  }
  size_t scratch_floats = 0;
  std::vector<int> fuse_at;
  for (size_t i = 0; split_on && i + 1 < pr->ops.size(); ++i) {
    StackOp& a = pr->ops[i];
    StackOp& b = pr->ops[i + 1];
    if (a.kind != kStGemm || b.kind != kStLayerNorm) continue;
    const StackGemmOp& g = a.u.g;
    const StackLnOp& l = b.u.ln;
    const bool last = i + 2 == pr->ops.size();
    if (g.out_split != nullptr || g.out_f32 == nullptr || g.f32_col0 != 0 || g.act != 0 || g.taps != 1 || l.x != g.out_f32 ||
        l.ld_x != g.ld_out || l.dim != g.n || a.rows != b.rows || a.rows_slot != b.rows_slot || !a.barrier_after ||
        !(b.barrier_after || last) || g.k_blocks < 8 || g.k_blocks % 4 != 0 || g.n % 32 != 0 || g.n > 512 || l.halo)
      continue;
    fuse_at.push_back((int)i);
    const int ksmax = (g.k_blocks % 8 == 0 && g.k_blocks >= 16) ? 8 : 4;
    const size_t need = (size_t)ksmax * a.rows * g.n;
    if (need > scratch_floats) scratch_floats = need;
  }
  if (!fuse_at.empty()) {
    for (int j = 0; j < 2; ++j)
      if (cudaMalloc(&pr->d_scratch[j], scratch_floats * sizeof(float)) != cudaSuccess) {
        set_error("stack_end: cudaMalloc of the split-K scratch (%zu bytes) failed", scratch_floats * sizeof(float));
        pr->failed = true;
        return FB200_ECUDA;
      }
    for (size_t f = 0; f < fuse_at.size(); ++f) {
      StackOp& a = pr->ops[fuse_at[f]];
      StackOp& b = pr->ops[fuse_at[f] + 1];
      StackGemmOp& g = a.u.g;
      StackLnOp& l = b.u.ln;
      float* scratch = pr->d_scratch[f & 1];
      const int ksmax = (g.k_blocks % 8 == 0 && g.k_blocks >= 16) ? 8 : 4;
      l.part = scratch;
      l.p_bias = g.bias;
      l.p_res = g.res1;
      l.ld_res = g.ld_res1;
      l.part_stride = (long long)a.rows * g.n;
      l.ld_part = g.n;
      l.g_ncols = g.n;
      l.g_kblocks = g.k_blocks;
      l.g_nparts = g.nparts;
      l.ksplit_max = ksmax;
      g.bias = nullptr;
      g.res1 = nullptr;
      g.out_f32 = scratch;
      g.ld_out = g.n;
      g.ksplit_max = ksmax;
      g.part_rows = a.rows;
      const int rc = make_tma_map(&pr->maps[g.map0], scratch, 4, (uint64_t)ksmax * a.rows, (uint64_t)g.n, (uint64_t)g.n, 32, 32, 128);
      if (rc) {
        pr->failed = true;
        return rc;
      }
    }
    pr->fused = (int)fuse_at.size();
  }
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
  unsigned long long* trace = pr->d_trace;
  if (trace != nullptr) cudaMemsetAsync(trace, 0, (pr->ops.size() * 8 + 8) * sizeof(unsigned long long), (cudaStream_t)stream);
  static int flags = -1;  // experiment switches (FB200_STACK_FLAGS): bit 0 = no L2 prefetch of the next op's weights / parameters
  if (flags < 0) {
    const char* e = getenv("FB200_STACK_FLAGS");
    flags = e ? atoi(e) : 0;
  }
  void* args[] = {(void*)&ops, (void*)&maps, (void*)&n_ops, (void*)&d, (void*)&sync_words, (void*)&trace, (void*)&flags};
  const cudaError_t e = cudaLaunchCooperativeKernel((const void*)stack_kernel, dim3((unsigned)pr->grid), dim3(kStackThreads), args,
                                                    kStSmemBytes, (cudaStream_t)stream);
  if (e != cudaSuccess) {
    set_error("stack_run: launch failed: %s", cudaGetErrorString(e));
    return FB200_ECUDA;
  }
  count_launch();
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_stack_profile(void* handle, int enable) {
  StackProgram* pr = (StackProgram*)handle;
  FB_CHECK_ARG(pr != nullptr && !pr->open && pr->d_ops != nullptr, "stack_profile: program not finished");
  if (enable && pr->d_trace == nullptr) {
    const size_t bytes = (pr->ops.size() * 8 + 8) * sizeof(unsigned long long);
    if (cudaMalloc(&pr->d_trace, bytes) != cudaSuccess || cudaMemset(pr->d_trace, 0, bytes) != cudaSuccess) {
      set_error("stack_profile: cudaMalloc failed");
      return FB200_ECUDA;
    }
  } else if (!enable && pr->d_trace != nullptr) {
    cudaDeviceSynchronize();
    cudaFree(pr->d_trace);
    pr->d_trace = nullptr;
  }
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_stack_profile_read(void* handle, int32_t* kinds, int32_t* barrier_after,
                                                                                double* stamp_us, int capacity, double* sm_mhz) {
  StackProgram* pr = (StackProgram*)handle;
  FB_CHECK_ARG(pr != nullptr && pr->d_trace != nullptr, "stack_profile_read: profiling is not enabled on this program");
  const int n = (int)pr->ops.size();
  FB_CHECK_ARG(capacity >= n, "stack_profile_read: capacity %d < %d ops", capacity, n);
  std::vector<unsigned long long> t((size_t)n * 8 + 8);
  if (cudaDeviceSynchronize() != cudaSuccess ||
      cudaMemcpy(t.data(), pr->d_trace, t.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost) != cudaSuccess) {
    set_error("stack_profile_read: device read failed");
    return FB200_ECUDA;
  }
  for (int i = 0; i < n; ++i) {
    if (kinds) kinds[i] = pr->ops[i].kind;
    if (barrier_after) barrier_after[i] = pr->ops[i].barrier_after;
    if (stamp_us)
      for (int k = 0; k < 8; ++k) stamp_us[i * 8 + k] = t[i * 8 + k] ? (double)(t[i * 8 + k] - t[0]) * 1e-3 : -1.0;
  }
  if (sm_mhz) {  // SM cycles between the first and the last stamp of CTA 0 over the wall time between them
    const double us = (double)(t[(size_t)(n - 1) * 8 + 7] - t[0]) * 1e-3;
    *sm_mhz = us > 0.0 ? (double)(t[(size_t)n * 8 + 1] - t[(size_t)n * 8]) / us : 0.0;
  }
  return n;
}

extern "C" __attribute__((visibility("default"))) int fb200_stack_destroy(void* handle) {
  StackProgram* pr = (StackProgram*)handle;
  if (pr == nullptr) return FB200_OK;
  if (g_rec == pr) g_rec = nullptr;
  if (pr->d_ops) cudaFree(pr->d_ops);
  if (pr->d_maps) cudaFree(pr->d_maps);
  if (pr->d_sync) cudaFree(pr->d_sync);
  if (pr->d_trace) cudaFree(pr->d_trace);
  for (int j = 0; j < 2; ++j)
    if (pr->d_scratch[j]) cudaFree(pr->d_scratch[j]);
  delete pr;
  return FB200_OK;
}
