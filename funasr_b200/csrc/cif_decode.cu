// CIF predictor tail (alpha, integrate-and-fire scan, frame integration) and greedy decoding
// (row argmax + log-prob, CTC collapse).  Integer/index work is bit-exact by construction.
#include <math.h>

#include "common.cuh"

namespace fb200 {

// One warp per packed row r; conv rows are in the halo layout m = r + seq(r).
__global__ void cif_alpha_kernel(const float* __restrict__ conv, int ld, const int32_t* __restrict__ cu_rows,
                                 int batch, int total_rows, int dim, const float* __restrict__ w,
                                 const float* __restrict__ bias, float smooth, float noise,
                                 float* __restrict__ alphas) {
  const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (r >= total_rows) return;
  const int m = r + find_seq(cu_rows, batch, r);
  const float* x = conv + (size_t)m * ld;
  float acc = 0.f;
  for (int c = lane; c < dim; c += 32) acc = fmaf(x[c], __ldg(w + c), acc);
  acc = warp_sum(acc);
  if (lane == 0) {
    const float z = acc + __ldg(bias);
    const float a = 1.f / (1.f + expf(-z));
    alphas[r] = fmaxf(a * smooth - noise, 0.f);
  }
}

// One warp per utterance: fp64 prefix sum as a warp scan over contiguous chunks.
__global__ void cif_scan_kernel(const float* __restrict__ alphas, const int32_t* __restrict__ cu_rows, int batch,
                                float tail, int32_t* __restrict__ token_num, int32_t* __restrict__ fire_pos,
                                float* __restrict__ fire_rem, float* __restrict__ fires) {
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (b >= batch) return;
  const int beg = cu_rows[b], t_len = cu_rows[b + 1] - beg;
  const int n = t_len + 1;  // appended tail frame
  const int chunk = (n + 31) / 32;
  const int lo = min(lane * chunk, n), hi = min(lo + chunk, n);
  const float* a = alphas + beg;
  double local = 0.0;
  for (int t = lo; t < hi; ++t) local += (double)(t < t_len ? a[t] : tail);
  double incl = local;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const double up = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl += up;
  }
  double run = incl - local;  // exclusive prefix of this lane's chunk
  const int obase = beg + b;  // per-utterance outputs have T_b + 1 slots
  float prev_floor = lo == 0 ? 0.f : floorf((float)run);
  for (int t = lo; t < hi; ++t) {
    run += (double)(t < t_len ? a[t] : tail);
    const float cs = (float)run;
    const float fl = floorf(cs);
    const bool fire = fl > prev_floor;
    const float fv = ((fire ? 1.f : 0.f) + cs) - fl;
    if (fires) fires[obase + t] = fv;
    if (fire) {
      const int kidx = (int)fl - 1;
      fire_pos[obase + kidx] = t;
      fire_rem[obase + kidx] = fv - floorf(fv);
    }
    prev_floor = fl;
  }
  const double total = __shfl_sync(0xffffffffu, incl, 31);
  if (lane == 0) token_num[b] = (int)floorf((float)total);
}

// One CTA per token; a thread owns 4 channels.
__global__ void cif_integrate_kernel(const float* __restrict__ hidden, int ld_h, const float* __restrict__ alphas,
                                     const int32_t* __restrict__ cu_rows, const int32_t* __restrict__ cu_tok, int batch,
                                     int dim, float tail, const int32_t* __restrict__ fire_pos,
                                     const float* __restrict__ fire_rem, float* __restrict__ embeds, int ld_e) {
  __shared__ int s_b;
  const int tok = blockIdx.x;
  if (threadIdx.x == 0) s_b = find_seq(cu_tok, batch, tok);
  __syncthreads();
  const int b = s_b;
  const int kidx = tok - cu_tok[b];
  const int beg = cu_rows[b], t_len = cu_rows[b + 1] - beg;
  const int obase = beg + b;
  const int t_k = fire_pos[obase + kidx];
  const int t_prev = kidx > 0 ? fire_pos[obase + kidx - 1] : -1;
  const float rem_k = fire_rem[obase + kidx];
  const float rem_prev = kidx > 0 ? fire_rem[obase + kidx - 1] : 0.f;
  for (int c = threadIdx.x * 4; c < dim; c += blockDim.x * 4) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int t = t_prev + 1; t <= t_k && t < t_len; ++t) {  // frame T_b is the appended zero row
      const float al = alphas[beg + t];
      const float4 h = *reinterpret_cast<const float4*>(hidden + (size_t)(beg + t) * ld_h + c);
      acc.x = fmaf(al, h.x, acc.x); acc.y = fmaf(al, h.y, acc.y);
      acc.z = fmaf(al, h.z, acc.z); acc.w = fmaf(al, h.w, acc.w);
    }
    if (t_prev >= 0 && t_prev < t_len) {
      const float4 h = *reinterpret_cast<const float4*>(hidden + (size_t)(beg + t_prev) * ld_h + c);
      acc.x = fmaf(rem_prev, h.x, acc.x); acc.y = fmaf(rem_prev, h.y, acc.y);
      acc.z = fmaf(rem_prev, h.z, acc.z); acc.w = fmaf(rem_prev, h.w, acc.w);
    }
    if (t_k < t_len) {
      const float4 h = *reinterpret_cast<const float4*>(hidden + (size_t)(beg + t_k) * ld_h + c);
      acc.x = fmaf(-rem_k, h.x, acc.x); acc.y = fmaf(-rem_k, h.y, acc.y);
      acc.z = fmaf(-rem_k, h.z, acc.z); acc.w = fmaf(-rem_k, h.w, acc.w);
    }
    *reinterpret_cast<float4*>(embeds + (size_t)tok * ld_e + c) = acc;
  }
}

// One CTA per row: argmax (first index on ties) and max log-prob = -log(sum exp(x - max)).
__global__ void __launch_bounds__(256)
argmax_logprob_kernel(const float* __restrict__ logits, int ld, int vocab, int32_t* __restrict__ ids,
                      float* __restrict__ score) {
  __shared__ float s_val[8];
  __shared__ int s_idx[8];
  __shared__ float s_sum[8];
  const float* x = logits + (size_t)blockIdx.x * ld;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int c = tid; c < vocab; c += 256) {
    const float vv = x[c];
    if (vv > best) { best = vv; bi = c; }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, off);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, off);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if (lane == 0) { s_val[warp] = best; s_idx[warp] = bi; }
  __syncthreads();
  best = s_val[0]; bi = s_idx[0];
#pragma unroll
  for (int wv = 1; wv < 8; ++wv)
    if (s_val[wv] > best || (s_val[wv] == best && s_idx[wv] < bi)) { best = s_val[wv]; bi = s_idx[wv]; }
  float sum = 0.f;
  for (int c = tid; c < vocab; c += 256) sum += expf(x[c] - best);
  sum = warp_sum(sum);
  if (lane == 0) s_sum[warp] = sum;
  __syncthreads();
  if (tid == 0) {
    float tot = 0.f;
#pragma unroll
    for (int wv = 0; wv < 8; ++wv) tot += s_sum[wv];
    ids[blockIdx.x] = bi;
    if (score) score[blockIdx.x] = -logf(tot);
  }
}

// One warp per utterance: neighbour compare + ballot compaction.
__global__ void ctc_collapse_kernel(const int32_t* __restrict__ ids, const int32_t* __restrict__ cu_rows, int batch,
                                    uint32_t drop_mask, int collapse, int32_t* __restrict__ out_ids,
                                    int32_t* __restrict__ out_len) {
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (b >= batch) return;
  const int beg = cu_rows[b], n = cu_rows[b + 1] - beg;
  int written = 0;
  for (int t0 = 0; t0 < n; t0 += 32) {
    const int t = t0 + lane;
    const int id = t < n ? ids[beg + t] : -1;
    const int prev = (t > 0 && t < n) ? ids[beg + t - 1] : -1;
    bool keep = t < n;
    if (collapse && t > 0 && id == prev) keep = false;
    if (id >= 0 && id < 32 && ((drop_mask >> id) & 1u)) keep = false;
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (keep) out_ids[beg + written + __popc(bal & ((1u << lane) - 1u))] = id;
    written += __popc(bal);
  }
  if (lane == 0) out_len[b] = written;
}

}  // namespace fb200

using namespace fb200;

extern "C" __attribute__((visibility("default"))) int fb200_cif_alpha(const float* conv_relu, int ld, const int32_t* cu_rows, int batch, int total_rows,
                               int dim, const float* w, const float* b, float smooth, float noise, float* alphas,
                               void* stream) {
  FB_CHECK_ARG(batch > 0 && dim > 0, "cif_alpha: bad batch/dim");
  if (total_rows == 0) return FB200_OK;
  ProfScope prof(kProfCifAlpha, stream, (double)total_rows * (dim + 1) * 4.0);
  cif_alpha_kernel<<<(total_rows + 7) / 8, 256, 0, (cudaStream_t)stream>>>(conv_relu, ld, cu_rows, batch, total_rows,
                                                                          dim, w, b, smooth, noise, alphas);
  FB_CHECK_LAUNCH("cif_alpha");
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_cif_scan(const float* alphas, const int32_t* cu_rows, int batch, float tail_threshold,
                              int32_t* token_num, int32_t* fire_pos, float* fire_rem, float* fires, void* stream) {
  FB_CHECK_ARG(batch > 0, "cif_scan: bad batch");
  ProfScope prof(kProfCifScan, stream, 0.0);
  cif_scan_kernel<<<(batch + 3) / 4, 128, 0, (cudaStream_t)stream>>>(alphas, cu_rows, batch, tail_threshold, token_num,
                                                                    fire_pos, fire_rem, fires);
  FB_CHECK_LAUNCH("cif_scan");
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_cif_integrate(const float* hidden, int ld_h, const float* alphas, const int32_t* cu_rows,
                                   const int32_t* cu_tok, int batch, int total_tok, int dim, float tail_threshold,
                                   const int32_t* fire_pos, const float* fire_rem, float* embeds, int ld_e,
                                   void* stream) {
  FB_CHECK_ARG(dim % 4 == 0 && ld_h % 4 == 0 && ld_e % 4 == 0, "cif_integrate: dim/ld must be multiples of 4");
  if (total_tok == 0) return FB200_OK;
  ProfScope prof(kProfCifIntegrate, stream, (double)total_tok * dim * 4.0);
  cif_integrate_kernel<<<total_tok, 128, 0, (cudaStream_t)stream>>>(hidden, ld_h, alphas, cu_rows, cu_tok, batch, dim,
                                                                   tail_threshold, fire_pos, fire_rem, embeds, ld_e);
  FB_CHECK_LAUNCH("cif_integrate");
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_argmax_logprob(const float* logits, int ld, int rows, int vocab, int32_t* ids, float* score,
                                    void* stream) {
  FB_CHECK_ARG(vocab > 0 && ld >= vocab, "argmax: bad vocab/ld");
  if (rows == 0) return FB200_OK;
  ProfScope prof(kProfArgmax, stream, (double)rows * vocab * 4.0);
  argmax_logprob_kernel<<<rows, 256, 0, (cudaStream_t)stream>>>(logits, ld, vocab, ids, score);
  FB_CHECK_LAUNCH("argmax_logprob");
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_ctc_collapse(const int32_t* ids, const int32_t* cu_rows, int batch, uint32_t drop_mask,
                                  int collapse_repeats, int32_t* out_ids, int32_t* out_len, void* stream) {
  FB_CHECK_ARG(batch > 0, "ctc_collapse: bad batch");
  ProfScope prof(kProfCtc, stream, 0.0);
  ctc_collapse_kernel<<<(batch + 3) / 4, 128, 0, (cudaStream_t)stream>>>(ids, cu_rows, batch, drop_mask,
                                                                        collapse_repeats, out_ids, out_len);
  FB_CHECK_LAUNCH("ctc_collapse");
  return FB200_OK;
}
