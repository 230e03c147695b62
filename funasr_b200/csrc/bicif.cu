// BiCifParaformer's time-stamp head on the device:
//   * fb200_lstm_bidir  -- the one-layer bidirectional LSTM of CifPredictorV3 (`self.blstm`,
//     funasr/models/bicif_paraformer/cif_predictor.py:142-145, :262-265) over a packed var-len batch, fused with the
//     dot product of `cif_output2` (:145, :270): a cooperative, persistent kernel, recurrent weights resident in shared
//     memory, one grid barrier per time step and direction;
//   * fb200_us_fires    -- sigmoid / smooth / noise / length scaling of the upsampled alphas (:270-283), the sequential
//     fp32 integrate-and-fire of `cif_wo_hidden` (:71-93) and the re-normalised second scan of
//     `ts_prediction_lfr6_standard` (funasr/utils/timestamp_tools.py:46-53), one warp per utterance.
#include <mutex>

#include "common.cuh"

namespace fb200 {
namespace lstm {

constexpr int H = 512;        // hidden units per direction
constexpr int UPC = 8;        // hidden units owned by one CTA (all four gates of each)
constexpr int CTAS = H / UPC; // CTAs per direction
constexpr int BC = 64;        // utterances staged per chunk
constexpr int HP = H + 4;     // shared-memory row pitch in floats (16-byte aligned, bank-shifted)
constexpr int THREADS = 256;
constexpr int MAX_CHUNKS = 16;
constexpr size_t kSmem = (size_t)(4 * UPC * HP + BC * HP + 128 * 16 + BC * UPC) * sizeof(float);

struct Params {
  const float* xg;      // [rows3, ld_xg]: W_ih x + b_ih + b_hh, columns [dir][gate i,f,g,o][H]
  int ld_xg;
  const int32_t* cu_rows;  // [batch + 1] prefix of the encoder rows; utterance b owns upsampled rows [up*cu[b], up*cu[b+1])
  int up, batch, rows3, max_len3;
  const float* w_hh;    // [2][4H][H]
  const float* w_out;   // [2][H]  cif_output2.weight split by direction
  float* hbuf;          // [2 dir][2 phase][batch][H]   zero on entry
  float* cbuf;          // [2 dir][batch][H]            zero on entry
  unsigned int* counters;  // [2]                        zero on entry
  float* part;          // [2 * CTAS][rows3] partial logits of every CTA
  float* out_h;         // optional [rows3][2H]
};

__device__ __forceinline__ unsigned int ld_acquire(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release(unsigned int* p) {
  asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p) : "memory");
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ float sigmoidf(float x) { return 1.f / (1.f + expf(-x)); }

__global__ void __launch_bounds__(THREADS, 1) lstm_bidir_kernel(const Params p) {
  extern __shared__ __align__(16) float smem[];
  float* Ws = smem;                    // [4 * UPC][HP]   row = gate * UPC + unit
  float* hs = Ws + 4 * UPC * HP;       // [BC][HP]
  float* red = hs + BC * HP;           // [16][128]
  float* lg = red + 128 * 16;          // [BC][UPC]
  __shared__ int chunk_max[MAX_CHUNKS];

  const int tid = threadIdx.x;
  const int dir = blockIdx.x / CTAS, cta = blockIdx.x % CTAS, u0 = cta * UPC;
  const int B = p.batch;
  const int nchunks = (B + BC - 1) / BC;
  for (int i = tid; i < 4 * UPC * (H / 4); i += THREADS) {
    const int row = i / (H / 4), k4 = i % (H / 4);
    const int gate = row / UPC, unit = row % UPC;
    const float4 v = __ldg(reinterpret_cast<const float4*>(p.w_hh + ((size_t)dir * 4 * H + gate * H + u0 + unit) * H) + k4);
    *reinterpret_cast<float4*>(Ws + row * HP + k4 * 4) = v;
  }
  if (tid < MAX_CHUNKS) {
    int m = 0;
    for (int b = tid * BC; b < min(B, (tid + 1) * BC); ++b) m = max(m, p.up * (p.cu_rows[b + 1] - p.cu_rows[b]));
    chunk_max[tid] = m;
  }
  __syncthreads();

  const int ks = tid >> 7, rem = tid & 127, u = rem >> 4, bg = rem & 15;
  const float wout = __ldg(p.w_out + dir * H + u0 + u);
  unsigned int* counter = p.counters + dir;

  for (int t = 0; t < p.max_len3; ++t) {
    const float* rd = p.hbuf + ((size_t)(dir * 2 + (t & 1)) * B) * H;
    float* wr = p.hbuf + ((size_t)(dir * 2 + ((t + 1) & 1)) * B) * H;
    if (t > 0) {
      if (tid == 0) {
        const unsigned int target = (unsigned int)t * CTAS;
        unsigned int spins = 0;
        while (ld_acquire(counter) < target) {
          if (++spins > (1u << 27)) {
            printf("funasr_b200 lstm: barrier timeout dir %d cta %d step %d\n", dir, cta, t);
            __trap();
          }
        }
      }
      __syncthreads();
    }
    for (int c = 0; c < nchunks; ++c) {
      if (t >= chunk_max[c]) continue;
      const int b0 = c * BC, nb = min(BC, B - b0);
      for (int i = tid; i < nb * (H / 4); i += THREADS) {
        const int row = i >> 7, k4 = i & 127;
        cp_async16(hs + row * HP + k4 * 4, rd + (size_t)(b0 + row) * H + k4 * 4);
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      // this thread's (utterance, unit) pairs: the input projections are fetched while the copy is in flight
      float xv[4][4];
      bool act[4];
      int rowi[4];
      if (ks == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int b = b0 + i * 16 + bg;
          act[i] = false;
          rowi[i] = 0;
          if (b < B) {
            const int beg = p.cu_rows[b], len3 = p.up * (p.cu_rows[b + 1] - beg);
            if (t < len3) {
              act[i] = true;
              rowi[i] = p.up * beg + (dir ? len3 - 1 - t : t);
            }
          }
#pragma unroll
          for (int g = 0; g < 4; ++g)
            xv[i][g] = act[i] ? __ldg(p.xg + (size_t)rowi[i] * p.ld_xg + dir * 4 * H + g * H + u0 + u) : 0.f;
        }
      }
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncthreads();
      float acc[4][4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int g = 0; g < 4; ++g) acc[i][g] = 0.f;
      const int kbeg = ks * (H / 2);
#pragma unroll 2
      for (int k = kbeg; k < kbeg + H / 2; k += 4) {
        float4 w[4];
#pragma unroll
        for (int g = 0; g < 4; ++g) w[g] = *reinterpret_cast<const float4*>(Ws + (g * UPC + u) * HP + k);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 hv = *reinterpret_cast<const float4*>(hs + (i * 16 + bg) * HP + k);
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            acc[i][g] = fmaf(hv.x, w[g].x, acc[i][g]);
            acc[i][g] = fmaf(hv.y, w[g].y, acc[i][g]);
            acc[i][g] = fmaf(hv.z, w[g].z, acc[i][g]);
            acc[i][g] = fmaf(hv.w, w[g].w, acc[i][g]);
          }
        }
      }
      if (ks == 1) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int g = 0; g < 4; ++g) red[(i * 4 + g) * 128 + rem] = acc[i][g];
      }
      __syncthreads();
      if (ks == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float logit = 0.f;
          if (act[i]) {
            const int b = b0 + i * 16 + bg;
            float gate[4];
#pragma unroll
            for (int g = 0; g < 4; ++g) gate[g] = acc[i][g] + red[(i * 4 + g) * 128 + rem] + xv[i][g];
            float* cp = p.cbuf + ((size_t)dir * B + b) * H + u0 + u;
            const float cn = sigmoidf(gate[1]) * *cp + sigmoidf(gate[0]) * tanhf(gate[2]);
            const float hn = sigmoidf(gate[3]) * tanhf(cn);
            *cp = cn;
            wr[(size_t)b * H + u0 + u] = hn;
            if (p.out_h) p.out_h[(size_t)rowi[i] * 2 * H + dir * H + u0 + u] = hn;
            logit = hn * wout;
          }
          lg[(i * 16 + bg) * UPC + u] = logit;
        }
      }
      __syncthreads();
      if (tid < nb) {
        const int b = b0 + tid;
        const int beg = p.cu_rows[b], len3 = p.up * (p.cu_rows[b + 1] - beg);
        if (t < len3) {
          float s = 0.f;
#pragma unroll
          for (int j = 0; j < UPC; ++j) s += lg[tid * UPC + j];
          p.part[(size_t)(dir * CTAS + cta) * p.rows3 + p.up * beg + (dir ? len3 - 1 - t : t)] = s;
        }
      }
      __syncthreads();
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) red_release(counter);
  }
}

// ---------------------------------------------------------------- tensor-core recurrence
// Same schedule as lstm_bidir_kernel with the h . W_hh^T product of a step on the tensor cores (warp-level
// mma.sync m16n8k16, bf16 operands split hi + mid, 3 terms, fp32 accumulate -- the operand scheme of the GEMMs): the SIMT
// kernel re-reads every staged h row once per hidden unit and is bound by shared-memory bandwidth (12.6 us per 64-utterance
// chunk and step); here the CTA's W_hh slice lives in REGISTERS as B fragments (warp = one gate x 8 units x one quarter
// of K, 8 k-steps x 2 parts; the quarters' partial sums meet in shared memory) and the staged h is read once per gate
// through ldmatrix.  h travels between CTAs as bf16 hi + mid.
constexpr int HPB = H + 8;  // bf16 row pitch: 1040 B, ldmatrix rows land in distinct banks
constexpr size_t kSmemMma = (size_t)2 * BC * HPB * sizeof(__nv_bfloat16) + (size_t)(4 * 4 * BC * UPC) * sizeof(float);

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void split_pair(float x, float y, uint32_t& hi, uint32_t& mid) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(x, y);
  const __nv_bfloat162 m = __floats2bfloat162_rn(x - __low2float(h), y - __high2float(h));
  hi = *reinterpret_cast<const uint32_t*>(&h);
  mid = *reinterpret_cast<const uint32_t*>(&m);
}

constexpr int MMA_THREADS = 512;  // 16 warps = 4 gates x 4 k-quarters

__global__ void __launch_bounds__(MMA_THREADS, 1) lstm_bidir_mma_kernel(const Params p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __nv_bfloat16* hs_hi = reinterpret_cast<__nv_bfloat16*>(smem_raw);  // [BC][HPB]
  __nv_bfloat16* hs_mid = hs_hi + BC * HPB;                           // [BC][HPB]
  float* gates_s = reinterpret_cast<float*>(hs_mid + BC * HPB);       // [4 k-quarters][4 gates][BC][UPC] partial sums
  __shared__ int chunk_max[MAX_CHUNKS];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int dir = blockIdx.x / CTAS, cta = blockIdx.x % CTAS, u0 = cta * UPC;
  const int B = p.batch;
  const int nchunks = (B + BC - 1) / BC;
  // this warp: gate nt (8 units) over k-steps [8 kq, 8 kq + 8) for all 64 staged rows -- four warps per scheduler, so the
  // ldmatrix latency of one warp hides under the products of the others
  const int nt = warp & 3, kq = warp >> 2;
  const int g = lane >> 2, t4 = lane & 3;
  uint32_t whi[8][2], wmid[8][2];  // B fragments of W_hh[dir][gate nt][unit u0 + g][128 kq ...], hi and mid
  {
    const float* wrow = p.w_hh + ((size_t)dir * 4 * H + nt * H + u0 + g) * H + 128 * kq;
#pragma unroll
    for (int s = 0; s < 8; ++s) {
      const float2 a = __ldg(reinterpret_cast<const float2*>(wrow + 16 * s + 2 * t4));
      const float2 b = __ldg(reinterpret_cast<const float2*>(wrow + 16 * s + 2 * t4 + 8));
      split_pair(a.x, a.y, whi[s][0], wmid[s][0]);
      split_pair(b.x, b.y, whi[s][1], wmid[s][1]);
    }
  }
  if (tid < MAX_CHUNKS) {
    int m = 0;
    for (int b = tid * BC; b < min(B, (tid + 1) * BC); ++b) m = max(m, p.up * (p.cu_rows[b + 1] - p.cu_rows[b]));
    chunk_max[tid] = m;
  }
  __syncthreads();

  __nv_bfloat16* hb = reinterpret_cast<__nv_bfloat16*>(p.hbuf);  // [dir][phase][B][hi 512 | mid 512]
  unsigned int* counter = p.counters + dir;
  const int eb = tid >> 3, eu = tid & 7;  // epilogue: utterance eb of the chunk, unit eu
  const float wout = __ldg(p.w_out + dir * H + u0 + eu);
  // ldmatrix row address of this lane inside a 16 x 16 tile: matrices (rows 0-7 | 8-15) x (k 0-7 | 8-15)
  const int lrow = (lane & 7) + ((lane >> 3) & 1) * 8, lcol = (lane >> 4) * 8 + 128 * kq;
  const uint32_t hi_base = (uint32_t)__cvta_generic_to_shared(hs_hi) + (uint32_t)(lrow * HPB + lcol) * 2u;
  const uint32_t mid_base = (uint32_t)__cvta_generic_to_shared(hs_mid) + (uint32_t)(lrow * HPB + lcol) * 2u;

  for (int t = 0; t < p.max_len3; ++t) {
    const __nv_bfloat16* rd = hb + ((size_t)(dir * 2 + (t & 1)) * B) * (2 * H);
    __nv_bfloat16* wr = hb + ((size_t)(dir * 2 + ((t + 1) & 1)) * B) * (2 * H);
    if (t > 0) {
      if (tid == 0) {
        const unsigned int target = (unsigned int)t * CTAS;
        unsigned int spins = 0;
        while (ld_acquire(counter) < target) {
          if (++spins > (1u << 27)) {
            printf("funasr_b200 lstm: barrier timeout dir %d cta %d step %d\n", dir, cta, t);
            __trap();
          }
        }
      }
      __syncthreads();
    }
    for (int c = 0; c < nchunks; ++c) {
      if (t >= chunk_max[c]) continue;
      const int b0 = c * BC, nb = min(BC, B - b0);
      for (int i = tid; i < nb * 128; i += MMA_THREADS) {
        const int row = i >> 7, c16 = i & 127, part = c16 >> 6, off = (c16 & 63) * 8;
        cp_async16((part ? hs_mid : hs_hi) + row * HPB + off, rd + (size_t)(b0 + row) * (2 * H) + part * H + off);
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      // the input projections of this thread's (utterance, unit) are fetched while the copy is in flight
      const int b = b0 + eb;
      bool act = false;
      int rowi = 0;
      if (b < B) {
        const int beg = p.cu_rows[b], len3 = p.up * (p.cu_rows[b + 1] - beg);
        if (t < len3) {
          act = true;
          rowi = p.up * beg + (dir ? len3 - 1 - t : t);
        }
      }
      float xv[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) xv[q] = act ? __ldg(p.xg + (size_t)rowi * p.ld_xg + dir * 4 * H + q * H + u0 + eu) : 0.f;
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncthreads();
      float acc[4][4];
#pragma unroll
      for (int mt = 0; mt < 4; ++mt)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[mt][q] = 0.f;
#pragma unroll
      for (int s = 0; s < 8; ++s) {
#pragma unroll
        for (int mt = 0; mt < 4; ++mt) {
          uint32_t ah[4], am[4];
          const uint32_t off = (uint32_t)(mt * 16 * HPB + s * 16) * 2u;
          ldmatrix_x4(ah, hi_base + off);
          ldmatrix_x4(am, mid_base + off);
          mma_bf16(acc[mt], ah, whi[s][0], whi[s][1]);
          mma_bf16(acc[mt], am, whi[s][0], whi[s][1]);
          mma_bf16(acc[mt], ah, wmid[s][0], wmid[s][1]);
        }
      }
      float* gq = gates_s + (size_t)(kq * 4 + nt) * BC * UPC;
#pragma unroll
      for (int mt = 0; mt < 4; ++mt) {
        const int row = mt * 16 + g;
        *reinterpret_cast<float2*>(gq + row * UPC + 2 * t4) = make_float2(acc[mt][0], acc[mt][1]);
        *reinterpret_cast<float2*>(gq + (row + 8) * UPC + 2 * t4) = make_float2(acc[mt][2], acc[mt][3]);
      }
      __syncthreads();
      {
        float logit = 0.f;
        if (act) {
          float gate[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float a = xv[q];
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) a += gates_s[((kk * 4 + q) * BC + eb) * UPC + eu];
            gate[q] = a;
          }
          float* cp = p.cbuf + ((size_t)dir * B + b) * H + u0 + eu;
          const float cn = sigmoidf(gate[1]) * *cp + sigmoidf(gate[0]) * tanhf(gate[2]);
          const float hn = sigmoidf(gate[3]) * tanhf(cn);
          *cp = cn;
          const __nv_bfloat16 hi = __float2bfloat16_rn(hn);
          wr[(size_t)b * (2 * H) + u0 + eu] = hi;
          wr[(size_t)b * (2 * H) + H + u0 + eu] = __float2bfloat16_rn(hn - __bfloat162float(hi));
          if (p.out_h) p.out_h[(size_t)rowi * 2 * H + dir * H + u0 + eu] = hn;
          logit = hn * wout;
        }
        // the CTA's share of the Linear(2H -> 1): the eight threads of an utterance hold its 8 units (fixed order)
        logit += __shfl_xor_sync(0xffffffffu, logit, 1);
        logit += __shfl_xor_sync(0xffffffffu, logit, 2);
        logit += __shfl_xor_sync(0xffffffffu, logit, 4);
        if (act && eu == 0) p.part[(size_t)(dir * CTAS + cta) * p.rows3 + rowi] = logit;
      }
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) red_release(counter);
  }
}

// ---------------------------------------------------------------- upsampled alphas -> fires
struct FireParams {
  const float* part;   // [nparts][rows3]
  int nparts;
  float bias, smooth, noise, threshold;
  const int32_t* cu_rows;
  int up, batch, rows3;
  const int32_t* token_num;  // [batch] predicted token counts (the length the alphas are scaled to)
  const int32_t* ids;        // packed token ids, utterance b owns [cu_tok[b], cu_tok[b+1])
  const int32_t* cu_tok;
  int drop0, drop1, drop2;   // ids not counted as characters (blank, sos, eos)
  float* us_alphas;          // [rows3]
  float* us_peaks;           // optional [rows3]: integrate value after each frame (first scan)
  int32_t* fire_idx;         // [rows3] fired frames of utterance b from up*cu[b], in order
  int32_t* fire_count;       // [batch]
  int32_t* num_chars;        // [batch]
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Sequential fp32 integrate-and-fire over alphas[0 .. n) * mul (or / div when div != 0): every lane carries the same
// scalar, lane i of a 32-frame chunk records the fire of frame i.  Returns the number of fires.
__device__ int fire_scan(const float* alphas, int n, float div, float thr, float* peaks, int32_t* fire_idx, int lane) {
  float integrate = 0.f;
  int count = 0;
  for (int base = 0; base < n; base += 32) {
    float mine = base + lane < n ? alphas[base + lane] : 0.f;
    if (div != 0.f) mine = mine / div;
    const int m = min(32, n - base);
    float my_peak = 0.f;
    bool my_fire = false;
    for (int i = 0; i < m; ++i) {
      const float a = __shfl_sync(0xffffffffu, mine, i);
      integrate = integrate + a;
      const bool fire = integrate >= thr;
      if (lane == i) {
        my_peak = integrate;
        my_fire = fire;
      }
      if (fire) integrate = integrate - thr;
    }
    if (peaks != nullptr && base + lane < n) peaks[base + lane] = my_peak;
    const unsigned int mask = __ballot_sync(0xffffffffu, my_fire);
    if (my_fire) fire_idx[count + __popc(mask & ((1u << lane) - 1u))] = base + lane;
    count += __popc(mask);
  }
  return count;
}

// alphas2 = relu(sigmoid(bias + sum of the LSTM kernel's partial dot products) * smooth - noise), one thread per frame
__global__ void __launch_bounds__(256) us_alpha_kernel(const FireParams p) {
  const int r = blockIdx.x * 256 + threadIdx.x;
  if (r >= p.rows3) return;
  float logit = p.bias;
  for (int j = 0; j < p.nparts; ++j) logit += __ldg(p.part + (size_t)j * p.rows3 + r);
  const float a = 1.f / (1.f + expf(-logit));
  p.us_alphas[r] = fmaxf(a * p.smooth - p.noise, 0.f);
}

__global__ void __launch_bounds__(128) us_fires_kernel(const FireParams p) {
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (b >= p.batch) return;
  const int beg3 = p.up * p.cu_rows[b], n = p.up * p.cu_rows[b + 1] - beg3;
  float* alphas = p.us_alphas + beg3;
  int32_t* fire_idx = p.fire_idx + beg3;
  float sum = 0.f;
  for (int r = lane; r < n; r += 32) sum += alphas[r];  // alphas2 from us_alpha_kernel
  sum = warp_sum(sum);
  const float scale = (float)p.token_num[b] / sum;
  float sum2 = 0.f;
  for (int r = lane; r < n; r += 32) {
    const float a = alphas[r] * scale;
    alphas[r] = a;
    sum2 += a;
  }
  sum2 = warp_sum(sum2);
  int chars = 0;
  for (int j = p.cu_tok[b] + lane; j < p.cu_tok[b + 1]; j += 32) {
    const int id = p.ids[j];
    chars += (id != p.drop0 && id != p.drop1 && id != p.drop2) ? 1 : 0;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) chars += __shfl_xor_sync(0xffffffffu, chars, o);
  __syncwarp();
  int count = fire_scan(alphas, n, 0.f, p.threshold, p.us_peaks ? p.us_peaks + beg3 : nullptr, fire_idx, lane);
  if (chars > 0 && count != chars + 1) {  // timestamp_tools.py:46-53: rescale so the alphas sum to chars + 1, fire again
    __syncwarp();
    count = fire_scan(alphas, n, sum2 / (float)(chars + 1), p.threshold, nullptr, fire_idx, lane);
  }
  if (lane == 0) {
    p.fire_count[b] = count;
    p.num_chars[b] = chars;
  }
}

}  // namespace lstm
}  // namespace fb200

using namespace fb200;

extern "C" __attribute__((visibility("default"))) int fb200_lstm_bidir(
    const float* xg, int ld_xg, const int32_t* cu_rows, int upsample, int batch, int rows3, int max_len3, int hidden,
    const float* w_hh, const float* w_out, float* workspace, long long workspace_bytes, float* part, float* out_h, int impl,
    void* stream) {
  using namespace fb200::lstm;
  FB_CHECK_ARG(impl == 0 || impl == 1, "lstm_bidir: impl must be 0 (tensor cores) or 1 (fp32 SIMT check kernel)");
  FB_CHECK_ARG(hidden == H, "lstm_bidir: hidden size %d unsupported (512 only)", hidden);
  FB_CHECK_ARG(batch > 0 && batch <= BC * MAX_CHUNKS, "lstm_bidir: batch %d outside 1..%d", batch, BC * MAX_CHUNKS);
  FB_CHECK_ARG(upsample >= 1 && ld_xg >= 8 * H, "lstm_bidir: upsample %d / ld_xg %d", upsample, ld_xg);
  if (rows3 == 0 || max_len3 == 0) return FB200_OK;
  const size_t h_bytes = (size_t)2 * 2 * batch * H * sizeof(float), c_bytes = (size_t)2 * batch * H * sizeof(float);
  const size_t need = h_bytes + c_bytes + 256;
  FB_CHECK_ARG(workspace != nullptr && (size_t)workspace_bytes >= need, "lstm_bidir: workspace %lld B < %zu B", workspace_bytes, need);
  cudaStream_t st = (cudaStream_t)stream;
  ProfScope prof(kProfLstm, stream, 2.0 * rows3 * 2.0 * 4.0 * H * H);
  cudaError_t e = cudaMemsetAsync(workspace, 0, need, st);
  if (e != cudaSuccess) {
    set_error("lstm_bidir: cudaMemsetAsync: %s", cudaGetErrorString(e));
    return FB200_ECUDA;
  }
  Params p;
  p.xg = xg;
  p.ld_xg = ld_xg;
  p.cu_rows = cu_rows;
  p.up = upsample;
  p.batch = batch;
  p.rows3 = rows3;
  p.max_len3 = max_len3;
  p.w_hh = w_hh;
  p.w_out = w_out;
  p.hbuf = workspace;
  p.cbuf = workspace + (size_t)2 * 2 * batch * H;
  p.counters = reinterpret_cast<unsigned int*>(workspace + (size_t)2 * 2 * batch * H + (size_t)2 * batch * H);
  p.part = part;
  p.out_h = out_h;
  static bool attr_set = false;
  if (!attr_set) {
    e = cudaFuncSetAttribute(lstm_bidir_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(lstm_bidir_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemMma);
    if (e != cudaSuccess) {
      set_error("lstm_bidir: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return FB200_ECUDA;
    }
    attr_set = true;
  }
  // every CTA waits for all the others of its direction at every step: they must be co-resident
  void* args[] = {(void*)&p};
  if (impl == 0)
    e = cudaLaunchCooperativeKernel((const void*)lstm_bidir_mma_kernel, dim3(2 * CTAS), dim3(MMA_THREADS), args, kSmemMma, st);
  else
    e = cudaLaunchCooperativeKernel((const void*)lstm_bidir_kernel, dim3(2 * CTAS), dim3(THREADS), args, kSmem, st);
  if (e != cudaSuccess) {
    set_error("lstm_bidir: cooperative launch: %s", cudaGetErrorString(e));
    return FB200_ECUDA;
  }
  count_launch();
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_lstm_workspace_bytes(int batch, int hidden, long long* bytes) {
  FB_CHECK_ARG(bytes != nullptr && batch > 0 && hidden > 0, "lstm_workspace_bytes: bad arguments");
  *bytes = (long long)((size_t)6 * batch * hidden * sizeof(float) + 256);
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_lstm_parts(int hidden, int* parts) {
  FB_CHECK_ARG(parts != nullptr && hidden == fb200::lstm::H, "lstm_parts: hidden size %d unsupported (512 only)", hidden);
  *parts = 2 * fb200::lstm::CTAS;
  return FB200_OK;
}

// ---------------------------------------------------------------- fb200_lstm_bidir without any allocation
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it): the
// LSTM's scratch (state double buffer, cell state, step counters) lives in a file-scope device array instead of a
// caller-provided workspace.  No counterpart in the reference.
//   * the array is sized for kStaticMaxBatch utterances (6 * 64 * 512 * 4 + 256 = 786 688 bytes); every GPU that loads
//     the library holds it until the process exits;
//   * in HOST code `g_lstm_static_scratch` is only a stand-in for the symbol, not a device address: the address is taken
//     once per device with cudaGetSymbolAddress (that device current) and handed to the kernel through Params;
//   * every call shares the array and finds the previous call's contents in it: the part the kernel needs at zero is
//     cleared by a cudaMemsetAsync on the caller's stream ahead of the launch.  Two calls that overlap on the GPU
//     (different streams, no dependency) race on it.
namespace fb200 {
namespace lstm {
constexpr int kStaticMaxBatch = 64;
constexpr size_t kStaticScratchBytes = (size_t)6 * kStaticMaxBatch * H * sizeof(float) + 256;
static __device__ __align__(256) unsigned char g_lstm_static_scratch[kStaticScratchBytes];

constexpr int kStaticMaxDevices = 64;
static float* s_static_ptr[kStaticMaxDevices] = {};  // per device: address of the array, set once under s_static_mu
static std::mutex s_static_mu;

// Per-device set-up with that device current: the symbol's address and the kernels' shared-memory opt-in.
static int static_scratch(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kStaticMaxDevices) {
    set_error("lstm_bidir_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return FB200_ECUDA;
  }
  std::lock_guard<std::mutex> lock(s_static_mu);
  float* ptr = s_static_ptr[dev];
  if (ptr == nullptr) {
    void* sym = nullptr;
    e = cudaGetSymbolAddress(&sym, g_lstm_static_scratch);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(lstm_bidir_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(lstm_bidir_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemMma);
    if (e != cudaSuccess || sym == nullptr) {
      set_error("lstm_bidir_static: set-up on device %d: %s", dev, cudaGetErrorString(e));
      return FB200_ECUDA;
    }
    ptr = s_static_ptr[dev] = static_cast<float*>(sym);
  }
  *out = ptr;
  return FB200_OK;
}
}  // namespace lstm
}  // namespace fb200

extern "C" __attribute__((visibility("default"))) int fb200_lstm_static_max_batch(void) { return fb200::lstm::kStaticMaxBatch; }

extern "C" __attribute__((visibility("default"))) int fb200_lstm_bidir_static(
    const float* xg, int ld_xg, const int32_t* cu_rows, int upsample, int batch, int rows3, int max_len3, int hidden,
    const float* w_hh, const float* w_out, float* part, float* out_h, int impl, void* stream) {
  using namespace fb200::lstm;
  FB_CHECK_ARG(impl == 0 || impl == 1, "lstm_bidir_static: impl must be 0 (tensor cores) or 1 (fp32 SIMT check kernel)");
  FB_CHECK_ARG(hidden == H, "lstm_bidir_static: hidden size %d unsupported (512 only)", hidden);
  FB_CHECK_ARG(batch > 0, "lstm_bidir_static: batch %d", batch);
  FB_CHECK_ARG(batch <= kStaticMaxBatch,
               "lstm_bidir_static: batch %d exceeds the static scratch (at most %d utterances, %zu bytes); use fb200_lstm_bidir "
               "with a caller-provided workspace",
               batch, kStaticMaxBatch, kStaticScratchBytes);
  FB_CHECK_ARG(upsample >= 1 && ld_xg >= 8 * H, "lstm_bidir_static: upsample %d / ld_xg %d", upsample, ld_xg);
  float* scratch = nullptr;
  const int rc = static_scratch(&scratch);
  if (rc != FB200_OK) return rc;
  if (rows3 == 0 || max_len3 == 0) return FB200_OK;  // set-up only (e.g. ahead of a graph capture)
  const size_t need = (size_t)6 * batch * H * sizeof(float) + 256;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(scratch, 0, need, st);  // the last call's state and counters are still in there
  if (e != cudaSuccess) {
    set_error("lstm_bidir_static: cudaMemsetAsync: %s", cudaGetErrorString(e));
    return FB200_ECUDA;
  }
  Params p;
  p.xg = xg;
  p.ld_xg = ld_xg;
  p.cu_rows = cu_rows;
  p.up = upsample;
  p.batch = batch;
  p.rows3 = rows3;
  p.max_len3 = max_len3;
  p.w_hh = w_hh;
  p.w_out = w_out;
  p.hbuf = scratch;
  p.cbuf = scratch + (size_t)2 * 2 * batch * H;
  p.counters = reinterpret_cast<unsigned int*>(scratch + (size_t)2 * 2 * batch * H + (size_t)2 * batch * H);
  p.part = part;
  p.out_h = out_h;
  void* args[] = {(void*)&p};
  if (impl == 0)
    e = cudaLaunchCooperativeKernel((const void*)lstm_bidir_mma_kernel, dim3(2 * CTAS), dim3(MMA_THREADS), args, kSmemMma, st);
  else
    e = cudaLaunchCooperativeKernel((const void*)lstm_bidir_kernel, dim3(2 * CTAS), dim3(THREADS), args, kSmem, st);
  if (e != cudaSuccess) {
    set_error("lstm_bidir_static: cooperative launch: %s", cudaGetErrorString(e));
    return FB200_ECUDA;
  }
  count_launch();
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_us_fires(
    const float* part, int nparts, int rows3, float bias, float smooth, float noise, float threshold, const int32_t* cu_rows,
    int upsample, int batch, const int32_t* token_num, const int32_t* ids, const int32_t* cu_tok, const int32_t* drop_ids,
    float* us_alphas, float* us_peaks, int32_t* fire_idx, int32_t* fire_count, int32_t* num_chars, void* stream) {
  FB_CHECK_ARG(batch > 0 && nparts >= 1 && upsample >= 1, "us_fires: batch %d / parts %d / upsample %d", batch, nparts, upsample);
  FB_CHECK_ARG(drop_ids != nullptr, "us_fires: drop_ids (3 host ints) missing");
  if (rows3 == 0) return FB200_OK;
  ProfScope prof(kProfUsFires, stream, (double)rows3 * (nparts + 4.0) * 4.0);
  fb200::lstm::FireParams p;
  p.part = part;
  p.nparts = nparts;
  p.bias = bias;
  p.smooth = smooth;
  p.noise = noise;
  p.threshold = threshold;
  p.cu_rows = cu_rows;
  p.up = upsample;
  p.batch = batch;
  p.rows3 = rows3;
  p.token_num = token_num;
  p.ids = ids;
  p.cu_tok = cu_tok;
  p.drop0 = drop_ids[0];
  p.drop1 = drop_ids[1];
  p.drop2 = drop_ids[2];
  p.us_alphas = us_alphas;
  p.us_peaks = us_peaks;
  p.fire_idx = fire_idx;
  p.fire_count = fire_count;
  p.num_chars = num_chars;
  fb200::lstm::us_alpha_kernel<<<(rows3 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(p);
  FB_CHECK_LAUNCH("us_alpha");
  fb200::lstm::us_fires_kernel<<<(batch + 3) / 4, 128, 0, (cudaStream_t)stream>>>(p);
  FB_CHECK_LAUNCH("us_fires");
  return FB200_OK;
}

// ---------------------------------------------------------------- SeACo: `_merge_res` on the greedy picks
// funasr/models/seaco_paraformer/model.py:312-322 with seaco_weight = 1: a token keeps the plain decoder's distribution when
// the hot-word distribution's argmax is NO_BIAS, else it takes the hot-word distribution; on the greedy path only the argmax
// and its log-probability of the chosen distribution are needed.
__global__ void merge_hotword_ids_kernel(const int32_t* __restrict__ dec_ids, const float* __restrict__ dec_logp,
                                         const int32_t* __restrict__ hw_ids, const float* __restrict__ hw_logp, int no_bias, int n,
                                         int32_t* __restrict__ out_ids, float* __restrict__ out_logp) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const bool keep = hw_ids[i] == no_bias;
  out_ids[i] = keep ? dec_ids[i] : hw_ids[i];
  if (out_logp != nullptr) out_logp[i] = keep ? dec_logp[i] : hw_logp[i];
}

extern "C" __attribute__((visibility("default"))) int fb200_merge_hotword_ids(
    const int32_t* dec_ids, const float* dec_logp, const int32_t* hw_ids, const float* hw_logp, int no_bias, int n, int32_t* out_ids,
    float* out_logp, void* stream) {
  FB_CHECK_ARG(n >= 0 && dec_ids && hw_ids && out_ids && (out_logp == nullptr || (dec_logp && hw_logp)), "merge_hotword_ids: bad arguments");
  if (n == 0) return FB200_OK;
  merge_hotword_ids_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(dec_ids, dec_logp, hw_ids, hw_logp, no_bias, n, out_ids, out_logp);
  FB_CHECK_LAUNCH("merge_hotword_ids");
  return FB200_OK;
}

// ---------------------------------------------------------------- SeACo: attention-score filter
// `hotword_scores` of `_seaco_decode_with_ASF` (funasr/models/seaco_paraformer/model.py:268-274): the cross-attention
// probabilities of SeACo decoder layer 5 (`get_attn_mat`, funasr/models/paraformer/decoder.py:123-137) summed over heads and
// tokens, per hot word.  One CTA per utterance walks (head, token) in a fixed order -- deterministic sums, the top-k that
// follows must not depend on scheduling.  fp32 on the CUDA cores: a few MFLOP per utterance.
namespace fb200 {
namespace lstm {
constexpr int kAsfThreads = 256;
constexpr int kAsfMaxWords = 2048;

__global__ void __launch_bounds__(kAsfThreads) asf_scores_kernel(const float* __restrict__ q, int ld_q, const float* __restrict__ k,
                                                                 int ld_k, const int32_t* __restrict__ cu_tok,
                                                                 const int32_t* __restrict__ cu_mem, int heads, int head_dim,
                                                                 float scale, float* __restrict__ scores) {
  __shared__ float s_q[128];
  __shared__ float s_p[kAsfMaxWords];
  __shared__ float s_acc[kAsfMaxWords];
  __shared__ float s_red[kAsfThreads / 32];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int t0 = cu_tok[b], nt = cu_tok[b + 1] - t0;
  const int m0 = cu_mem[b], nm = cu_mem[b + 1] - m0;
  for (int n = tid; n < nm; n += kAsfThreads) s_acc[n] = 0.f;
  __syncthreads();
  for (int h = 0; h < heads; ++h) {
    for (int l = 0; l < nt; ++l) {
      if (tid < head_dim) s_q[tid] = q[(size_t)(t0 + l) * ld_q + h * head_dim + tid] * scale;
      __syncthreads();
      float mx = -INFINITY;
      for (int n = tid; n < nm; n += kAsfThreads) {
        const float* kr = k + (size_t)(m0 + n) * ld_k + h * head_dim;
        float d = 0.f;
        for (int c = 0; c < head_dim; c += 4) {
          const float4 kv = *reinterpret_cast<const float4*>(kr + c);
          d = fmaf(s_q[c], kv.x, d);
          d = fmaf(s_q[c + 1], kv.y, d);
          d = fmaf(s_q[c + 2], kv.z, d);
          d = fmaf(s_q[c + 3], kv.w, d);
        }
        s_p[n] = d;
        mx = fmaxf(mx, d);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      if (lane == 0) s_red[warp] = mx;
      __syncthreads();
      mx = s_red[0];
      for (int w = 1; w < kAsfThreads / 32; ++w) mx = fmaxf(mx, s_red[w]);
      __syncthreads();
      float sum = 0.f;
      for (int n = tid; n < nm; n += kAsfThreads) {
        const float e = expf(s_p[n] - mx);
        s_p[n] = e;
        sum += e;
      }
      sum = warp_sum(sum);
      if (lane == 0) s_red[warp] = sum;
      __syncthreads();
      sum = 0.f;
      for (int w = 0; w < kAsfThreads / 32; ++w) sum += s_red[w];
      const float inv = 1.f / sum;
      for (int n = tid; n < nm; n += kAsfThreads) s_acc[n] += s_p[n] * inv;
      __syncthreads();
    }
  }
  for (int n = tid; n < nm; n += kAsfThreads) scores[m0 + n] = s_acc[n];
}
}  // namespace lstm
}  // namespace fb200

extern "C" __attribute__((visibility("default"))) int fb200_asf_scores(
    const float* q, int ld_q, const float* k, int ld_k, const int32_t* cu_tok, const int32_t* cu_mem, int batch, int max_words, int heads,
    int head_dim, float scale, float* scores, void* stream) {
  FB_CHECK_ARG(batch > 0 && heads > 0 && head_dim > 0 && head_dim <= 128 && head_dim % 4 == 0 && ld_q % 4 == 0 && ld_k % 4 == 0,
               "asf_scores: batch %d / heads %d / head_dim %d", batch, heads, head_dim);
  FB_CHECK_ARG(max_words >= 1 && max_words <= fb200::lstm::kAsfMaxWords, "asf_scores: %d hot words per utterance (1..%d)", max_words,
               fb200::lstm::kAsfMaxWords);
  fb200::lstm::asf_scores_kernel<<<batch, fb200::lstm::kAsfThreads, 0, (cudaStream_t)stream>>>(q, ld_q, k, ld_k, cu_tok, cu_mem, heads, head_dim,
                                                                                             scale, scores);
  FB_CHECK_LAUNCH("asf_scores");
  return FB200_OK;
}
