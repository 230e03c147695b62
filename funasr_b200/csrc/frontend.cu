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

// ---------------------------------------------------------------------------------------------------------------
// nfft = 512: one warp turns TWO frames into log-mel energies with ONE complex 512-point FFT (frame A in the real
// part, frame B in the imaginary part; the spectra are separated afterwards through X[k] and conj(X[N-k])).
// The FFT is 8 x 8 x 8: three passes of in-register radix-8 butterflies (two per lane and pass) with two exchanges
// through shared memory, instead of nine radix-2 passes through shared memory.
//   n = p + 64 j,  k = q + 8 (r + 8 k''),  p = p0 + 8 p1:
//   pass 1: y[p][q]      = W512^(p q)  * sum_j  x[p + 64 j]        W8^(j q)
//   pass 2: z[q][p0][r]  = W64^(p0 r)  * sum_p1 y[p0 + 8 p1][q]    W8^(p1 r)
//   pass 3: X[q+8r+64k''] =              sum_p0 z[q][p0][r]        W8^(p0 k'')
constexpr int kF5Warps = 4;
constexpr int kF5Ex = 576;     // exchange buffer entries: 8 rows of 72 (pass 1 -> 2), 8 of 65 (2 -> 3), 512 (spectrum)
constexpr int kF5Pw = 264;     // power spectrum entries per frame (257 used)
constexpr int kF5MelW = 1024;  // compacted non-zero mel weights (80 triangular filters over 257 bins: ~520)

// In-place forward DFT of 8 complex values, natural order in and out (decimation in frequency).
__device__ __forceinline__ void dft8(float (&re)[8], float (&im)[8]) {
  constexpr float h = 0.70710678118654752440f;
  float br[8], bi[8];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    br[j] = re[j] + re[j + 4];
    bi[j] = im[j] + im[j + 4];
    br[j + 4] = re[j] - re[j + 4];
    bi[j + 4] = im[j] - im[j + 4];
  }
  {  // W8^1 = h (1 - i), W8^2 = -i, W8^3 = -h (1 + i)
    const float a5 = br[5], b5 = bi[5], a6 = br[6], b6 = bi[6], a7 = br[7], b7 = bi[7];
    br[5] = h * (a5 + b5);
    bi[5] = h * (b5 - a5);
    br[6] = b6;
    bi[6] = -a6;
    br[7] = h * (b7 - a7);
    bi[7] = -h * (a7 + b7);
  }
#pragma unroll
  for (int g = 0; g < 2; ++g) {  // 4-point DFT of the even-output half (g = 0) and the odd-output half (g = 1)
    const int o = g * 4;
    const float d0r = br[o] + br[o + 2], d0i = bi[o] + bi[o + 2];
    const float d2r = br[o] - br[o + 2], d2i = bi[o] - bi[o + 2];
    const float d1r = br[o + 1] + br[o + 3], d1i = bi[o + 1] + bi[o + 3];
    const float d3r = bi[o + 1] - bi[o + 3], d3i = br[o + 3] - br[o + 1];  // (c1 - c3) * (-i)
    re[g + 0] = d0r + d1r;
    im[g + 0] = d0i + d1i;
    re[g + 2] = d2r + d3r;
    im[g + 2] = d2i + d3i;
    re[g + 4] = d0r - d1r;
    im[g + 4] = d0i - d1i;
    re[g + 6] = d2r - d3r;
    im[g + 6] = d2i - d3i;
  }
}

__global__ void __launch_bounds__(kF5Warps * 32)
fbank512_kernel(const float* __restrict__ wave, const int64_t* __restrict__ cu_samples,
                const int32_t* __restrict__ cu_frames, int batch, int total_frames,
                const float* __restrict__ window, int frame_len, int frame_shift,
                const float* __restrict__ mel_bank, int n_mels, float upscale, float preemph,
                float* __restrict__ out) {
  constexpr int N = 512, NB = N / 2 + 1;
  __shared__ float s_re[kF5Warps][kF5Ex];
  __shared__ float s_im[kF5Warps][kF5Ex];
  __shared__ float s_pw[kF5Warps][2][kF5Pw];
  __shared__ float2 s_tw[N];  // W512^t = (cos, -sin)(2 pi t / 512)
  __shared__ float s_win[N];
  __shared__ float s_melw[kF5MelW];
  __shared__ int s_lo[kMaxMel], s_hi[kMaxMel], s_off[kMaxMel];
  __shared__ int s_compact;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int t = tid; t < N; t += blockDim.x) {
    float sn, cs;
    sincospif(-2.0f * (float)t / (float)N, &sn, &cs);
    s_tw[t] = make_float2(cs, sn);
    s_win[t] = t < frame_len ? __ldg(window + t) : 0.f;
  }
  // non-zero span of every mel filter: one warp per filter
  for (int m = warp; m < n_mels; m += kF5Warps) {
    const float* row = mel_bank + (size_t)m * NB;
    int lo = NB, hi = 0;
    for (int k = lane; k < NB; k += 32)
      if (__ldg(row + k) != 0.0f) {
        lo = min(lo, k);
        hi = max(hi, k + 1);
      }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
      hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if (lane == 0) {
      s_lo[m] = hi > lo ? lo : 0;
      s_hi[m] = hi > lo ? hi : 0;
    }
  }
  __syncthreads();
  if (tid == 0) {
    int tot = 0;
    for (int m = 0; m < n_mels; ++m) {
      s_off[m] = tot;
      tot += s_hi[m] - s_lo[m];
    }
    s_compact = tot <= kF5MelW;
  }
  __syncthreads();
  const bool compact = s_compact != 0;
  if (compact)
    for (int m = warp; m < n_mels; m += kF5Warps)
      for (int k = s_lo[m] + lane; k < s_hi[m]; k += 32) s_melw[s_off[m] + k - s_lo[m]] = __ldg(mel_bank + (size_t)m * NB + k);
  __syncthreads();

  float* ex_re = s_re[warp];
  float* ex_im = s_im[warp];
  const int npairs = (total_frames + 1) / 2;
  for (int pr = blockIdx.x * kF5Warps + warp; pr < npairs; pr += gridDim.x * kF5Warps) {
    // ---- load both frames: remove DC, pre-emphasis, window.  Element t of a lane is sample n = lane + 32 t.
    float v[2][16];
#pragma unroll
    for (int f = 0; f < 2; ++f) {
      const int frame = 2 * pr + f;
      if (frame < total_frames) {
        const int b = find_seq(cu_frames, batch, frame);
        const float* x = wave + cu_samples[b] + (int64_t)(frame - cu_frames[b]) * frame_shift;
        float acc = 0.f;
#pragma unroll
        for (int t = 0; t < 16; ++t) {
          const int n = lane + 32 * t;
          v[f][t] = n < frame_len ? __ldg(x + n) * upscale : 0.f;
          acc += v[f][t];
        }
        const float mean = warp_sum(acc) / (float)frame_len;
#pragma unroll
        for (int t = 0; t < 16; ++t) {
          const int n = lane + 32 * t;
          float y = 0.f;
          if (n < frame_len) {
            const float prv = __ldg(x + (n > 0 ? n - 1 : 0)) * upscale - mean;
            y = ((v[f][t] - mean) - preemph * prv) * s_win[n];
          }
          v[f][t] = y;
        }
      } else {
#pragma unroll
        for (int t = 0; t < 16; ++t) v[f][t] = 0.f;
      }
    }
    // ---- pass 1: lane owns p = lane + 32 u, all j (sample p + 64 j is element t = u + 2 j)
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      float re[8], im[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        re[j] = v[0][u + 2 * j];
        im[j] = v[1][u + 2 * j];
      }
      dft8(re, im);
      const int pp = lane + 32 * u;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float2 w = s_tw[(pp * q) & (N - 1)];
        ex_re[q * 72 + pp] = re[q] * w.x - im[q] * w.y;
        ex_im[q * 72 + pp] = re[q] * w.y + im[q] * w.x;
      }
    }
    __syncwarp();
    // ---- pass 2: lane owns (q, p0) = (c >> 3, c & 7), c = lane + 32 u, all p1
    {
      float re[2][8], im[2][8];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int c = lane + 32 * u, q = c >> 3, p0 = c & 7;
#pragma unroll
        for (int p1 = 0; p1 < 8; ++p1) {
          re[u][p1] = ex_re[q * 72 + p0 + 8 * p1];
          im[u][p1] = ex_im[q * 72 + p0 + 8 * p1];
        }
      }
      __syncwarp();
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int c = lane + 32 * u, q = c >> 3, p0 = c & 7;
        dft8(re[u], im[u]);
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          const float2 w = s_tw[(8 * p0 * r) & (N - 1)];
          ex_re[p0 * 65 + 8 * q + r] = re[u][r] * w.x - im[u][r] * w.y;
          ex_im[p0 * 65 + 8 * q + r] = re[u][r] * w.y + im[u][r] * w.x;
        }
      }
    }
    __syncwarp();
    // ---- pass 3: lane owns (q, r) = (c >> 3, c & 7), all p0; the spectrum goes back in natural order
    {
      float re[2][8], im[2][8];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int c = lane + 32 * u;
#pragma unroll
        for (int p0 = 0; p0 < 8; ++p0) {
          re[u][p0] = ex_re[p0 * 65 + c];
          im[u][p0] = ex_im[p0 * 65 + c];
        }
      }
      __syncwarp();
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int c = lane + 32 * u, q = c >> 3, r = c & 7;
        dft8(re[u], im[u]);
#pragma unroll
        for (int k2 = 0; k2 < 8; ++k2) {
          ex_re[q + 8 * r + 64 * k2] = re[u][k2];
          ex_im[q + 8 * r + 64 * k2] = im[u][k2];
        }
      }
    }
    __syncwarp();
    // ---- separate the two real spectra and square: A = (Z[k] + conj Z[N-k]) / 2, B = (Z[k] - conj Z[N-k]) / (2 i)
#pragma unroll
    for (int i = 0; i < 9; ++i) {
      const int k = lane + 32 * i;
      if (k < NB) {
        const float a = ex_re[k], b = ex_im[k];
        const float c = ex_re[(N - k) & (N - 1)], d = ex_im[(N - k) & (N - 1)];
        const float ar = 0.5f * (a + c), ai = 0.5f * (b - d);
        const float br = 0.5f * (b + d), bi = 0.5f * (c - a);
        s_pw[warp][0][k] = ar * ar + ai * ai;
        s_pw[warp][1][k] = br * br + bi * bi;
      }
    }
    __syncwarp();
    // ---- mel filters + log
    for (int idx = lane; idx < 2 * n_mels; idx += 32) {
      const int f = idx >= n_mels ? 1 : 0, m = idx - f * n_mels;
      const int frame = 2 * pr + f;
      const int lo = s_lo[m], hi = s_hi[m];
      const float* pw = s_pw[warp][f];
      float e = 0.f;
      if (compact) {
        const float* wm = s_melw + s_off[m] - lo;
        for (int k = lo; k < hi; ++k) e = fmaf(pw[k], wm[k], e);
      } else {
        const float* row = mel_bank + (size_t)m * NB;
        for (int k = lo; k < hi; ++k) e = fmaf(pw[k], __ldg(row + k), e);
      }
      if (frame < total_frames) out[(size_t)frame * n_mels + m] = logf(fmaxf(e, FLT_EPSILON));
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
  if (nfft == 512) {
    const int npairs = (total_frames + 1) / 2;
    int g5 = (npairs + kF5Warps - 1) / kF5Warps;
    if (g5 > 148 * 5) g5 = 148 * 5;  // 38 KB of shared memory per CTA: five CTAs per SM, persistent
    fbank512_kernel<<<g5, kF5Warps * 32, 0, (cudaStream_t)stream>>>(wave, cu_samples, cu_frames, batch, total_frames, window,
                                                                    frame_len, frame_shift, mel_bank, n_mels, upscale, preemph,
                                                                    out);
  } else {
    fbank_kernel<<<grid, kFbankWarps * 32, 0, (cudaStream_t)stream>>>(
        wave, cu_samples, cu_frames, batch, total_frames, window, frame_len, frame_shift, nfft, log2n,
        mel_bank, n_mels, upscale, preemph, out);
  }
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
