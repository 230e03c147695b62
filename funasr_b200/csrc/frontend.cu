// Frontend kernels: Kaldi fbank (one warp per frame, 512-point FFT in shared memory) and
// LFR stacking + CMVN (+ encoder prologue).  HBM-bound gather/FMA work; see DESIGN.md.
#include <float.h>

#include "common.cuh"

namespace fb200 {

constexpr int kFbankWarps = 4;
constexpr int kMaxFft = 512;
constexpr int kMaxMel = 128;

// One warp turns one frame into n_mels log-mel energies.
//   smem per warp: re[nfft], im[nfft]; per CTA: twiddles[nfft/2] (cos, -sin), mel ranges.
__global__ void __launch_bounds__(kFbankWarps * 32)
fbank_kernel(const float* __restrict__ wave, const int64_t* __restrict__ cu_samples,
             const int32_t* __restrict__ cu_frames, int batch, int total_frames,
             const float* __restrict__ window, int frame_len, int frame_shift, int nfft, int log2n,
             const float* __restrict__ mel_bank, int n_mels, float upscale, float preemph,
             float* __restrict__ out) {
  __shared__ float s_re[kFbankWarps][kMaxFft];
  __shared__ float s_im[kFbankWarps][kMaxFft];
  __shared__ float s_twc[kMaxFft / 2];
  __shared__ float s_tws[kMaxFft / 2];
  __shared__ int s_lo[kMaxMel];
  __shared__ int s_hi[kMaxMel];

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int nbins = nfft / 2 + 1;

  for (int k = tid; k < nfft / 2; k += blockDim.x) {
    float s, c;
    sincospif(-2.0f * (float)k / (float)nfft, &s, &c);
    s_twc[k] = c;
    s_tws[k] = s;
  }
  // Non-zero span of every triangular mel filter (found once per CTA; CTAs are persistent).
  for (int m = tid; m < n_mels; m += blockDim.x) {
    const float* row = mel_bank + (size_t)m * nbins;
    int lo = nbins, hi = 0;
    for (int k = 0; k < nbins; ++k) {
      if (__ldg(row + k) != 0.0f) {
        if (k < lo) lo = k;
        hi = k + 1;
      }
    }
    s_lo[m] = lo;
    s_hi[m] = hi;
  }
  __syncthreads();

  float* re = s_re[warp];
  float* im = s_im[warp];
  const int shift_bits = 32 - log2n;

  for (int frame = blockIdx.x * kFbankWarps + warp; frame < total_frames; frame += gridDim.x * kFbankWarps) {
    const int b = find_seq(cu_frames, batch, frame);
    const float* x = wave + cu_samples[b] + (int64_t)(frame - cu_frames[b]) * frame_shift;

    // mean (remove_dc_offset)
    float acc = 0.f;
    for (int j = lane; j < frame_len; j += 32) acc += __ldg(x + j) * upscale;
    const float mean = warp_sum(acc) / (float)frame_len;

    // pre-emphasis + window, scattered to bit-reversed slots for the in-place DIT FFT
    for (int j = lane; j < nfft; j += 32) {
      float v = 0.f;
      if (j < frame_len) {
        const float cur = __ldg(x + j) * upscale - mean;
        const float prv = __ldg(x + (j > 0 ? j - 1 : 0)) * upscale - mean;
        v = (cur - preemph * prv) * __ldg(window + j);
      }
      const int r = (int)(__brev((unsigned)j) >> shift_bits);
      re[r] = v;
      im[r] = 0.f;
    }
    __syncwarp();

    for (int s = 0; s < log2n; ++s) {
      const int half = 1 << s;
      const int tw_step = (nfft >> 1) >> s;
      for (int i = lane; i < nfft / 2; i += 32) {
        const int pos = i & (half - 1);
        const int a = ((i >> s) << (s + 1)) + pos;
        const int bidx = a + half;
        const float wc = s_twc[pos * tw_step], ws = s_tws[pos * tw_step];
        const float br = re[bidx], bi = im[bidx];
        const float tr = br * wc - bi * ws;
        const float ti = br * ws + bi * wc;
        const float ar = re[a], ai = im[a];
        re[a] = ar + tr;
        im[a] = ai + ti;
        re[bidx] = ar - tr;
        im[bidx] = ai - ti;
      }
      __syncwarp();
    }

    // power spectrum into re[0..nbins)
    for (int k = lane; k < nbins; k += 32) {
      const float a = re[k], c = im[k];
      im[k] = a * a + c * c;
    }
    __syncwarp();

    for (int m = lane; m < n_mels; m += 32) {
      const float* row = mel_bank + (size_t)m * nbins;
      float e = 0.f;
      for (int k = s_lo[m]; k < s_hi[m]; ++k) e = fmaf(im[k], __ldg(row + k), e);
      out[(size_t)frame * n_mels + m] = logf(fmaxf(e, FLT_EPSILON));
    }
    __syncwarp();
  }
}

// One CTA per output row; float4 per thread.
__global__ void lfr_cmvn_kernel(const float* __restrict__ fbank, const int32_t* __restrict__ cu_frames,
                                const int32_t* __restrict__ cu_rows, int batch, int n_mels, int lfr_m,
                                int lfr_n, const float* __restrict__ shift, const float* __restrict__ scale,
                                const float* __restrict__ inv_ts, float pe_scale, int row_skip,
                                float* __restrict__ out, int ld_out) {
  __shared__ int s_b;
  const int row = blockIdx.x;
  if (threadIdx.x == 0) s_b = find_seq(cu_rows, batch, row);
  __syncthreads();
  const int b = s_b;
  const int i = row - cu_rows[b] - row_skip;
  if (i < 0) return;
  const int f0 = cu_frames[b], nf = cu_frames[b + 1] - f0;
  const int dim = n_mels * lfr_m;
  const int half = dim / 2;
  for (int c = threadIdx.x * 4; c < dim; c += blockDim.x * 4) {
    const int j = c / n_mels, cc = c - j * n_mels;
    int src = lfr_n * i - (lfr_m - 1) / 2 + j;
    src = min(max(src, 0), nf - 1);
    float4 v = *reinterpret_cast<const float4*>(fbank + (size_t)(f0 + src) * n_mels + cc);
    float r[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float y = r[u];
      if (shift != nullptr) y = (y + __ldg(shift + c + u)) * __ldg(scale + c + u);
      if (inv_ts != nullptr) {
        const int col = c + u;
        const float arg = (float)(i + 1 + row_skip) * __ldg(inv_ts + (col < half ? col : col - half));
        y = y * pe_scale + (col < half ? sinf(arg) : cosf(arg));
      }
      r[u] = y;
    }
    *reinterpret_cast<float4*>(out + (size_t)row * ld_out + c) = make_float4(r[0], r[1], r[2], r[3]);
  }
}

__global__ void scale_add_pe_kernel(const float* __restrict__ x, int ld_x, const int32_t* __restrict__ cu_rows,
                                    int batch, int dim, const float* __restrict__ inv_ts, float scale,
                                    float* __restrict__ out, int ld_out) {
  __shared__ int s_b;
  const int row = blockIdx.x;
  if (threadIdx.x == 0) s_b = find_seq(cu_rows, batch, row);
  __syncthreads();
  const int pos = row - cu_rows[s_b] + 1;
  const int half = dim / 2;
  for (int c = threadIdx.x; c < dim; c += blockDim.x) {
    const float arg = (float)pos * __ldg(inv_ts + (c < half ? c : c - half));
    out[(size_t)row * ld_out + c] = x[(size_t)row * ld_x + c] * scale + (c < half ? sinf(arg) : cosf(arg));
  }
}

// out[cu_rows[b] + j, :] = head[j, :] * scale + PE(pos = j + 1) for the first n_head rows of every sequence.
__global__ void fill_head_rows_kernel(const float* __restrict__ head, int n_head, int dim, const int32_t* __restrict__ cu_rows,
                                      const float* __restrict__ inv_ts, float scale, float* __restrict__ out, int ld_out) {
  const int b = blockIdx.x / n_head, j = blockIdx.x % n_head;
  const size_t row = (size_t)cu_rows[b] + j;
  const int half = dim / 2;
  for (int c = threadIdx.x; c < dim; c += blockDim.x) {
    float v = head[(size_t)j * dim + c] * scale;
    if (inv_ts != nullptr) {
      const float arg = (float)(j + 1) * __ldg(inv_ts + (c < half ? c : c - half));
      v += c < half ? sinf(arg) : cosf(arg);
    }
    out[row * ld_out + c] = v;
  }
}

}  // namespace fb200

using namespace fb200;

extern "C" __attribute__((visibility("default"))) int fb200_fbank(const float* wave, const int64_t* cu_samples, const int32_t* cu_frames, int batch,
                           int total_frames, const float* window, int frame_len, int frame_shift, int nfft,
                           const float* mel_bank, int n_mels, float upscale, float preemph, float* out,
                           void* stream) {
  FB_CHECK_ARG(batch > 0 && total_frames >= 0, "fbank: bad batch/frames");
  FB_CHECK_ARG(nfft >= 64 && nfft <= kMaxFft && (nfft & (nfft - 1)) == 0, "fbank: nfft %d unsupported", nfft);
  FB_CHECK_ARG(frame_len > 0 && frame_len <= nfft, "fbank: frame_len %d > nfft %d", frame_len, nfft);
  FB_CHECK_ARG(n_mels > 0 && n_mels <= kMaxMel, "fbank: n_mels %d unsupported", n_mels);
  if (total_frames == 0) return FB200_OK;
  ProfScope prof(kProfFbank, stream, 4.0 * (double)total_frames * (frame_shift + n_mels));
  int log2n = 0;
  while ((1 << log2n) < nfft) ++log2n;
  int grid = (total_frames + kFbankWarps - 1) / kFbankWarps;
  const int cap = 148 * 8;  // persistent: a multiple of the SM count
  if (grid > cap) grid = cap;
  fbank_kernel<<<grid, kFbankWarps * 32, 0, (cudaStream_t)stream>>>(
      wave, cu_samples, cu_frames, batch, total_frames, window, frame_len, frame_shift, nfft, log2n,
      mel_bank, n_mels, upscale, preemph, out);
  FB_CHECK_LAUNCH("fbank");
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_lfr_cmvn(const float* fbank, const int32_t* cu_frames, const int32_t* cu_rows, int batch,
                              int total_rows, int n_mels, int lfr_m, int lfr_n, const float* cmvn_shift,
                              const float* cmvn_scale, const float* pe_inv_timescales, float pe_scale,
                              int row_skip, float* out, int ld_out, void* stream) {
  FB_CHECK_ARG(n_mels % 4 == 0 && ld_out % 4 == 0, "lfr_cmvn: n_mels and ld_out must be multiples of 4");
  FB_CHECK_ARG((cmvn_shift == nullptr) == (cmvn_scale == nullptr), "lfr_cmvn: shift/scale must come together");
  FB_CHECK_ARG(ld_out >= n_mels * lfr_m, "lfr_cmvn: ld_out too small");
  if (total_rows == 0) return FB200_OK;
  ProfScope prof(kProfLfr, stream, 4.0 * (double)total_rows * n_mels * (lfr_n + lfr_m));
  lfr_cmvn_kernel<<<total_rows, 160, 0, (cudaStream_t)stream>>>(fbank, cu_frames, cu_rows, batch, n_mels, lfr_m,
                                                               lfr_n, cmvn_shift, cmvn_scale, pe_inv_timescales,
                                                               pe_scale, row_skip, out, ld_out);
  FB_CHECK_LAUNCH("lfr_cmvn");
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_scale_add_pe(const float* x, int ld_x, const int32_t* cu_rows, int batch, int total_rows,
                                  int dim, const float* pe_inv_timescales, float scale, float* out, int ld_out,
                                  void* stream) {
  FB_CHECK_ARG(dim % 2 == 0, "scale_add_pe: dim must be even");
  if (total_rows == 0) return FB200_OK;
  ProfScope prof(kProfPe, stream, 8.0 * (double)total_rows * dim);
  scale_add_pe_kernel<<<total_rows, 256, 0, (cudaStream_t)stream>>>(x, ld_x, cu_rows, batch, dim,
                                                                   pe_inv_timescales, scale, out, ld_out);
  FB_CHECK_LAUNCH("scale_add_pe");
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_fill_head_rows(const float* head, int n_head, int dim,
                                                                           const int32_t* cu_rows, int batch,
                                                                           const float* pe_inv_timescales, float scale,
                                                                           float* out, int ld_out, void* stream) {
  FB_CHECK_ARG(n_head > 0 && dim > 0 && dim % 2 == 0 && batch >= 0, "fill_head_rows: n_head %d dim %d batch %d", n_head, dim, batch);
  if (batch == 0) return FB200_OK;
  ProfScope prof(kProfPe, stream, 8.0 * (double)batch * n_head * dim);
  fill_head_rows_kernel<<<batch * n_head, 256, 0, (cudaStream_t)stream>>>(head, n_head, dim, cu_rows, pe_inv_timescales, scale,
                                                                         out, ld_out);
  FB_CHECK_LAUNCH("fill_head_rows");
  return FB200_OK;
}
