/*
 * funasr_b200 -- C-ABI of the B200-native (sm_100a) kernels behind FunASR's batched offline
 * non-autoregressive recognition path (WavFrontend -> SANMEncoder -> CifPredictorV2 ->
 * ParaformerSANMDecoder -> greedy / CTC greedy).
 *
 * Boundary rules:
 *   - plain pointers and sizes only; every pointer is DEVICE memory unless named `h_`;
 *   - the caller owns every buffer (inputs, outputs, workspaces); the library never allocates
 *     or frees device memory and keeps no mutable global state but a thread-local error string;
 *   - all work is enqueued on the caller's `stream` (a cudaStream_t passed as void*); no call
 *     synchronises the device;
 *   - return 0 on success, a negative FB200_E* code otherwise; text via fb200_last_error().
 *
 * Data layout: utterances are PACKED, never padded.  Row r of a packed matrix belongs to the
 * utterance b with cu_rows[b] <= r < cu_rows[b+1] (int32 prefix sums, cu_rows[0] = 0).  This is
 * the reference's batch_size=1 semantics (funasr/auto/auto_model.py:185-187) for a whole batch.
 *
 * "Split operand" (dense contractions): a GEMM operand X[rows, K] is stored as bf16
 * [rows, nparts * Kp], Kp = K rounded up to 64, part p in columns [p*Kp, (p+1)*Kp):
 *   nparts = 1: bf16(X)                       (throughput mode)
 *   nparts = 2: X ~= hi + mid, each bf16      (fp32-parity mode "bf16x3": 3 cross terms on tcgen05)
 *   nparts = 3: X ~= hi + mid + lo            (6 cross terms; check mode)
 * Padding columns are zero.
 *
 * Each entry point names the reference code it replaces (paths relative to the reference root).
 */
#ifndef FUNASR_B200_H_
#define FUNASR_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FB200_OK 0
#define FB200_EINVAL (-1)    /* bad shape / argument */
#define FB200_ECUDA (-2)     /* CUDA runtime or driver error */
#define FB200_EUNSUPPORTED (-3)

#define FB200_VERSION 100

int fb200_version(void);
const char* fb200_last_error(void);
/* Number of kernels launched by this thread through the library since the last reset. */
long long fb200_launch_count(int reset);

/* Optional per-entry-point device timing for the calling thread: while enabled every call below
 * is bracketed by CUDA events on its stream.  `kind` indexes the entry points in header order:
 * 0 fbank, 1 lfr_cmvn, 2 scale_add_pe, 3 layernorm, 4 split_rows/split_weight, 5 fsmn, 6 gemm,
 * 7 attention, 8 cif_alpha, 9 cif_scan, 10 cif_integrate, 11 argmax_logprob, 12 ctc_collapse, 13 conv2d_sub1,
 * 14 glu_dwconv, 15 attention_relpos, 16 lstm_bidir, 17 us_fires (gemm_ex / gemm_heads count as gemm, gather_rows as
 * split_rows).
 * fb200_profile_read synchronises on the recorded events and returns the summed milliseconds,
 * call count and algorithmic work (bytes for HBM-bound kinds, FLOPs for gemm/attention). */
#define FB200_PROFILE_KINDS 18
int fb200_profile_enable(int on);
int fb200_profile_read(int kind, double* ms, long long* calls, double* work, int reset);

/* ------------------------------------------------------------------ frontend --------- */

/* Kaldi fbank, replaces torchaudio.compliance.kaldi.fbank as called by WavFrontend.forward
 * (funasr/frontends/wav_frontend.py:127-143): frames of `frame_len` samples every `frame_shift`,
 * snip_edges, x32768 upscale (`upscale`), DC removal, pre-emphasis 0.97, window table,
 * zero-pad to `nfft`, |rfft|^2, mel filterbank [n_mels, nfft/2+1], log(max(., FLT_EPSILON)).
 *   wave        f32 [total_samples]   packed waveforms
 *   cu_samples  i64 [batch+1]         sample offsets
 *   cu_frames   i32 [batch+1]         frame offsets (frames_b = 1 + (N_b - frame_len)/frame_shift)
 *   window      f32 [frame_len]
 *   mel_bank    f32 [n_mels, nfft/2+1]
 *   out         f32 [total_frames, n_mels]
 */
int fb200_fbank(const float* wave, const int64_t* cu_samples, const int32_t* cu_frames, int batch,
                int total_frames, const float* window, int frame_len, int frame_shift, int nfft,
                const float* mel_bank, int n_mels, float upscale, float preemph, float* out,
                void* stream);

/* LFR stacking + CMVN (+ optional encoder prologue), replaces apply_lfr / apply_cmvn
 * (funasr/frontends/wav_frontend.py:41-74) and, when `pe_inv_timescales` != NULL, the
 * `xs_pad * sqrt(d)` + SinusoidalPositionEncoder prologue of SANMEncoder.forward
 * (funasr/models/sanm/encoder.py:378,397; funasr/models/transformer/embedding.py:383-413).
 *   out[row, n_mels*j + c] = (fbank[clamp(lfr_n*i - (lfr_m-1)/2 + j, 0, F-1), c] + shift) * scale
 *   cmvn_shift / cmvn_scale  f32 [n_mels*lfr_m] or NULL
 *   pe_inv_timescales        f32 [n_mels*lfr_m/2] or NULL; position = i + 1 + pos_offset
 *   out f32 [total_rows, ld_out] (ld_out >= n_mels*lfr_m), row = cu_rows[b] + row_offset_in_seq + i
 * `row_skip` rows at the head of every sequence are left untouched (SenseVoice prompt frames).
 */
int fb200_lfr_cmvn(const float* fbank, const int32_t* cu_frames, const int32_t* cu_rows, int batch,
                   int total_rows, int n_mels, int lfr_m, int lfr_n, const float* cmvn_shift,
                   const float* cmvn_scale, const float* pe_inv_timescales, float pe_scale,
                   int row_skip, float* out, int ld_out, void* stream);

/* Prompt rows of SenseVoiceSmall (funasr/models/sense_voice/model.py:852-876 prepends 4 embedding rows to
 * the features of every utterance; SenseVoiceEncoderSmall.forward :561-563 then scales and adds the position
 * code): out[cu_rows[b] + j, :] = head[j, :] * scale + PE(pos = j + 1), j < n_head, for every sequence b.
 * Used with fb200_lfr_cmvn(row_skip = n_head), which writes the speech rows behind them.
 *   head f32 [n_head, dim]; pe_inv_timescales f32 [dim/2] or NULL (no position code) */
int fb200_fill_head_rows(const float* head, int n_head, int dim, const int32_t* cu_rows, int batch,
                         const float* pe_inv_timescales, float scale, float* out, int ld_out,
                         void* stream);

/* Encoder prologue alone for callers that hand over features: out = x*scale + PE(pos = i+1). */
int fb200_scale_add_pe(const float* x, int ld_x, const int32_t* cu_rows, int batch, int total_rows,
                       int dim, const float* pe_inv_timescales, float scale, float* out, int ld_out,
                       void* stream);

/* ------------------------------------------------------------------ row-wise ops ------ */

/* LayerNorm over the last dim, replaces funasr/models/transformer/layer_norm.py:13-39
 * (eps 1e-12) and the SenseVoice copy (eps 1e-5).  Outputs (either may be NULL):
 *   out_f32   f32  [rows, ld_f32]
 *   out_split bf16 split operand [rows_out, nparts*Kp]; row r is written at
 *             r + (halo ? 1 + seq(r) : 0): `halo` interleaves one untouched (zero) row before
 *             every sequence and after the last, for the k=3 CIF convolution; the kernel
 *             writes those rows - 1 + batch of them - as zeros (sequences must be non-empty),
 *             so rows_out = rows + batch + 1.
 */
int fb200_layernorm(const float* x, int ld_x, int rows, int dim, const float* gamma,
                    const float* beta, float eps, float* out_f32, int ld_f32, void* out_split,
                    int nparts, int halo, const int32_t* cu_rows, int batch, void* stream);

/* fp32 [rows, dim] -> bf16 split operand (no normalisation). */
int fb200_split_rows(const float* x, int ld_x, int rows, int dim, void* out_split, int nparts,
                     void* stream);
/* Same, into the halo layout of fb200_layernorm (rows + batch + 1 output rows, interleaved zero
 * rows written by the kernel): the input of the CIF conv for callers that hold encoder output. */
int fb200_split_rows_halo(const float* x, int ld_x, int rows, int dim, void* out_split, int nparts,
                          const int32_t* cu_rows, int batch, void* stream);

/* Layout adapters for callers that hold the reference's padded tensors (pad_sequence layout,
 * funasr/frontends/wav_frontend.py:157): padded f32 [batch, max_len, dim] <-> packed [sum len, dim].
 * pack: row t of utterance b is read at padded + b*batch_stride + t*row_stride (elements).
 * unpack: writes a dense [batch, max_len, dim]; rows t >= len_b are filled with pad_value. */
int fb200_pack_rows(const float* padded, long long batch_stride, int row_stride, const int32_t* cu_rows,
                    int batch, int total_rows, int dim, float* packed, int ld_packed, void* stream);
int fb200_unpack_rows(const float* packed, int ld_packed, const int32_t* cu_rows, int batch, int max_len,
                      int dim, float pad_value, float* padded, void* stream);

/* FSMN memory block: depthwise conv1d over time, kernel `ksize`, zero halo PER SEQUENCE, plus
 * identity: out = conv(v) + v (+ residual).  Replaces MultiHeadedAttentionSANM.forward_fsmn
 * (funasr/models/sanm/attention.py:207-223) and MultiHeadedAttentionSANMDecoder.forward
 * (:499-547).  weight f32 [dim, ksize] (= Conv1d weight [dim,1,ksize]); left pad (ksize-1)/2.
 */
int fb200_fsmn(const float* v, int ld_v, const int32_t* cu_rows, int batch, int total_rows, int dim,
               const float* weight, int ksize, const float* residual, int ld_res, float* out,
               int ld_out, void* stream);

/* ------------------------------------------------------------------ dense contractions - */

/* C = A . W^T over split operands on tcgen05 (TMA -> smem -> UMMA, fp32 accumulators in TMEM).
 * Replaces every nn.Linear on the path and the k=3 Conv1d of CifPredictorV2:
 *   linear_q_k_v / linear_out (funasr/models/sanm/attention.py:225-252,290),
 *   w_1 / w_2 (funasr/models/transformer/positionwise_feed_forward.py:14-34,
 *              funasr/models/sanm/positionwise_feed_forward.py:12-33),
 *   linear_q / linear_k_v (attention.py:631-662), output_layer (paraformer/decoder.py:404),
 *   cif_conv1d (paraformer/cif_predictor.py:190,213), ctc_lo (funasr/models/ctc/ctc.py:173-197).
 *   a        bf16 split [a_rows, a_parts*Kp]
 *   w        bf16 split [n, w_parts*Kp] (per tap: [taps][n, w_parts*Kp], tap stride = n*w_parts*Kp)
 *   taps     1 for a Linear; 3 for the CIF conv: tap j reads A row (m + j) (halo layout)
 *   terms    cross terms summed: a_parts=w_parts=1 -> 1; 3x3 parts -> 6 (hi.hi, hi.mid, mid.hi,
 *            hi.lo, lo.hi, mid.mid) accumulated small-to-large
 *   epilogue y = acc + bias[n]; if relu y = max(y,0); y += res1[m, :]
 *   outputs  out_f32 [m, ld_out] holding columns [f32_col0, n) (column c at c - f32_col0; f32_col0 % 32 == 0), and/or
 *            out_split (bf16 split, out_parts parts, Kp = n up to 64)
 * impl: 0 = tcgen05 (product path), 1 = SIMT fp32 check kernel (same operands, for bring-up tests).
 */
int fb200_gemm(const void* a, int a_rows, int a_parts, const void* w, int w_parts, int m, int n,
               int k, int taps, const float* bias, int relu, const float* res1, int ld_res1,
               float* out_f32, int ld_out, int f32_col0, void* out_split, int out_parts, int impl,
               void* stream);

/* fb200_gemm with a general A operand, for contractions whose rows are gathered by fixed offsets:
 *   a            bf16, row pitch `lda` elements (0: a_parts*Kp), part p of the K columns at
 *                a_col0 + tap_col[j] + p * a_part_stride (0: Kp)
 *   h_tap_row    HOST int32 [taps]: tap j of output row m reads A row m + h_tap_row[j]; rows outside [0, a_rows)
 *                read as zeros.  NULL: tap j reads row m + j.   h_tap_col: HOST int32 [taps] or NULL (zeros).
 *   w            [taps][n, w_parts*Kp] as for fb200_gemm; taps <= 24
 *   act          0 none, 1 ReLU, 2 swish y*sigmoid(y) (funasr/models/transformer/utils/nets_utils.py:496-528)
 * Uses on the Conformer path (BASELINE config 4):
 *   - second 3x3 stride-2 convolution of Conv2dSubsampling (funasr/models/transformer/utils/subsampling.py:66) as a
 *     9-tap contraction over the parity planes written by fb200_conv2d_sub1 (tap (dt,df): row offset
 *     ((dt&1)*2 + (df&1)) * plane_rows + (dt>>1)*FP + (df>>1));
 *   - its Linear over (channel, frequency) pairs (:70,91) as a 19-tap contraction with COLUMN offsets over the
 *     conv output viewed as [segment rows, FP * nparts * C];
 *   - (q + pos_bias_v) . p^T per head (funasr/models/transformer/attention.py:303-305): A = one head's 64 columns of
 *     the fused q_u | q_v | k | v projection (a_col0, a_part_stride);
 *   - FFN w_1 with swish (funasr/models/transformer/positionwise_feed_forward.py:14-34). */
int fb200_gemm_ex(const void* a, int a_rows, int lda, int a_part_stride, int a_col0, int a_parts, const void* w,
                  int w_parts, int m, int n, int k, int taps, const int32_t* h_tap_row, const int32_t* h_tap_col,
                  const float* bias, int act, const float* res1, int ld_res1, float* out_f32, int ld_out, int f32_col0,
                  void* out_split, int out_parts, int impl, void* stream);

/* `groups` independent products over the SAME rows in one launch (the per-head (q+v).p^T of the rel-pos attention):
 *   out[r, g*P + c] = sum_k A[r, a_col0 + g*a_col_step + k] * W[g*w_group_rows + c, k],  c < P = n_group rounded up to 256
 * (columns n_group <= c < P of a group hold products with the weight rows that follow -- callers ignore them).
 * a as for fb200_gemm_ex; w bf16 split [groups * w_group_rows, w_parts*Kp]; out f32 [m, ld_out >= groups*P]. */
int fb200_gemm_heads(const void* a, int a_rows, int lda, int a_part_stride, int a_col0, int a_col_step, int a_parts,
                     const void* w, int w_group_rows, int w_parts, int m, int n_group, int groups, int k, float* out_f32,
                     int ld_out, int impl, void* stream);

/* fp32 weight [n, k] -> bf16 split [n, nparts*Kp] (done once at load). */
int fb200_split_weight(const float* w, int n, int k, void* out_split, int nparts, void* stream);

/* ------------------------------------------------------------------ attention ---------- */

/* Work list of fb200_attention for one (cu_q, cu_kv) pair: every 128-row query block of every
 * utterance that has keys.  Built on the device (one small kernel), reused by every attention call
 * over the same batch (all layers).  plan: int32 [4 + 4 * capacity]; capacity >= q_rows / 128 + batch
 * always suffices.  plan[0] = entries written, plan[1] = entries needed; a plan cut short (plan[1] > plan[0]) makes
 * fb200_attention / fb200_attention_relpos print the two numbers and trap (a sticky CUDA error on the stream) instead of
 * silently skipping query blocks. */
int fb200_attention_plan(const int32_t* cu_q, const int32_t* cu_kv, int batch, int32_t* plan,
                         int capacity, void* stream);

/* softmax(q k^T * scale) v per utterance and head, var-len, streaming softmax (scores never
 * materialised) on tcgen05: a persistent grid walks the plan's (query block, head) items;
 * S = Q K^T and O = P V as UMMAs with Q and P in tensor memory and TMA-staged K / V tiles, fp32
 * accumulators in TMEM.  Replaces MultiHeadedAttentionSANM.forward_attention
 * (funasr/models/sanm/attention.py:254-311) when cu_q == cu_kv, and MultiHeadedAttentionCrossAtt
 * (attention.py:664-717) otherwise.
 *   q, k, v   bf16 split matrices (as written by fb200_gemm's out_split): [q_rows, ld_q] and
 *             [kv_rows, ld_k / ld_v]; head h of part p starts at column
 *             {q,k,v}_col0 + p * {q,kv}_part_stride + h * head_dim.  k and v may be one matrix.
 *   nparts    1 (bf16) or 2 (hi + mid: 3 cross terms per contraction); 3 runs the SIMT kernel
 *   plan      from fb200_attention_plan(cu_q, cu_kv, ...); required by impl 0, ignored by impl 1
 *   out       f32 [q_rows, ld_out] and/or bf16 split [q_rows, out_parts * Kp], Kp = heads*head_dim
 * impl: 0 = tcgen05 (product path), 1 = SIMT fp32 check kernel over the same operands.
 */
int fb200_attention(const void* q, int ld_q, int q_col0, int q_part_stride, const void* k, int ld_k,
                    int k_col0, const void* v, int ld_v, int v_col0, int kv_part_stride, int nparts,
                    const int32_t* cu_q, const int32_t* cu_kv, const int32_t* plan, int batch,
                    int q_rows, int kv_rows, int max_q, int heads, int head_dim, float scale,
                    float* out_f32, int ld_out, void* out_split, int out_parts, int impl,
                    void* stream);

/* ------------------------------------------------------------------ Conformer ---------- */

/* First convolution of Conv2dSubsampling (funasr/models/transformer/utils/subsampling.py:64-65: Conv2d(1, C, 3, 2)
 * + ReLU) over packed fbank frames, written as a bf16 split operand in four parity planes so that the second
 * convolution is a 9-tap fb200_gemm_ex.  Per utterance b: F_b = cu_frames[b+1]-cu_frames[b] frames,
 * T1 = (F_b-3)/2+1, T2 = (T1-3)/2+1, a segment of T2+1 rows (cu_seg = prefix sums of T2_b + 1);
 * F1 = (n_mels-3)/2+1, FP = (F1+1)/2 slots per segment row, plane_rows = total_seg * FP.
 *   planes  bf16 [4 * plane_rows, nparts*C]; plane (pt,pf) row (seg_b + t')*FP + f' = relu(conv)(t1=2t'+pt, f1=2f'+pf),
 *           zeros where t1 >= T1 or f1 >= F1 (every slot is written)
 *   weight  f32 [C, 9] (= Conv2d weight [C,1,3,3]), bias f32 [C]; C even, <= 256 */
int fb200_conv2d_sub1(const float* feats, int ld_f, const int32_t* cu_frames, const int32_t* cu_seg, int batch,
                      int total_seg, int n_mels, const float* weight, const float* bias, int channels, void* planes,
                      int nparts, void* stream);

/* dst[cu_dst[b] + t, :] = src[cu_src[b] + t, :] for t < cu_dst[b+1] - cu_dst[b]: drops the scratch row each
 * utterance's segment carries after the subsampling contractions. */
int fb200_gather_rows(const float* src, int ld_src, const int32_t* cu_src, float* dst, int ld_dst,
                      const int32_t* cu_dst, int batch, int rows, int dim, void* stream);

/* ConvolutionModule between its pointwise convolutions (funasr/models/conformer/encoder.py:110-116):
 * GLU over g = [a | gate] -> depthwise Conv1d(C, C, ksize, padding (ksize-1)/2, groups=C) with bias, zero halo PER
 * UTTERANCE -> BatchNorm1d in eval mode folded by the caller into bn_scale = gamma/sqrt(var+eps),
 * bn_shift = beta - mean*bn_scale -> swish.  g f32 [rows, ld_g >= 2C]; weight f32 [C, ksize]; ksize 7, 15 or 31;
 * out_split bf16 split [rows, nparts*Kp(C)] (the A operand of pointwise_conv2). */
int fb200_glu_dwconv(const float* g, int ld_g, const int32_t* cu_rows, int batch, int total_rows, int max_len,
                     int channels, const float* weight, int ksize, const float* bias, const float* bn_scale,
                     const float* bn_shift, void* out_split, int nparts, void* stream);

/* LegacyRelPositionMultiHeadedAttention core (funasr/models/transformer/attention.py:271-312 with rel_shift
 * :249-269 and forward_attention :69-102), var-len, per utterance and head, streaming softmax:
 *   s[i,j] = ((q_i+u).k_j + shift(bd)[i,j]) * scale;  shift(bd)[i,j] = bd[i, T-1-i+j] (j<=i), 0 (j=i+1),
 *   bd[i+1, j-i-2] (j>=i+2)  -- the reference's pad/view/slice in closed form (T = utterance length)
 *   q (already q+u), k, v  bf16 split matrices, head h of part p at col0 + p*part_stride + h*head_dim
 *   bd  f32 [heads][rows, ld_bd] (head stride bd_head_stride elements): bd[r, c] = (q_r+v) . p_c, c < T of r's utterance
 *   out f32 [rows, ld_out] and/or bf16 split [rows, out_parts*Kp(heads*head_dim)];  head_dim 64.
 *   plan from fb200_attention_plan(cu_rows, cu_rows, ...) (128-row query blocks); required by impl 0.
 * impl: 0 = tcgen05 (S = Q K^T and P V as UMMAs, P in tensor memory, the shifted bias band staged through shared
 * memory; product path), 1 = SIMT fp32 check kernel over the same operands (also the 3-part mode). */
int fb200_attention_relpos(const void* q, int ld_q, int q_col0, int q_part_stride, const void* k, int ld_k, int k_col0,
                           const void* v, int ld_v, int v_col0, int kv_part_stride, int nparts, const float* bd,
                           int ld_bd, long long bd_head_stride, const int32_t* cu_rows, const int32_t* plan, int batch,
                           int rows, int max_len, int heads, int head_dim, float scale, float* out_f32, int ld_out,
                           void* out_split, int out_parts, int impl, void* stream);

/* ------------------------------------------------------------------ CIF predictor ------ */

/* alpha = sigmoid(relu_conv[m, :] . w + b) * smooth - noise, clamped at 0; rows are in the halo
 * layout of the conv GEMM (row m = r + seq(r)); writes alphas in PACKED order [total_rows].
 * Replaces cif_output + sigmoid (funasr/models/paraformer/cif_predictor.py:216-219). */
int fb200_cif_alpha(const float* conv_relu, int ld, const int32_t* cu_rows, int batch, int total_rows,
                    int dim, const float* w, const float* b, float smooth, float noise, float* alphas,
                    void* stream);

/* Integrate-and-fire scan per utterance: tail_threshold appended at index T_b
 * (tail_process_fn, cif_predictor.py:346-370), prefix sum in fp64 rounded to fp32, fire where
 * floor crosses (cif_wo_hidden_v1, :668-693).
 *   token_num  i32 [batch]    floor(sum alphas incl. tail)
 *   fire_pos   i32 [total_rows + batch] packed per utterance at cu_rows[b] + b: frame index of the
 *              k-th fire (index T_b = the appended zero frame)
 *   fire_rem   f32 same layout: frac(fires) at the fire (`remains`, cif_v1 :718)
 *   fires      f32 [total_rows + batch] or NULL: the reference's `fires`/cif_peak row (T_b+1 each)
 */
int fb200_cif_scan(const float* alphas, const int32_t* cu_rows, int batch, float tail_threshold,
                   int32_t* token_num, int32_t* fire_pos, float* fire_rem, float* fires, void* stream);

/* frames: embeds[cu_tok[b] + k, :] = sum_{t in (t_{k-1}, t_k]} alpha_t h_t + rem_{k-1} h_{t_{k-1}}
 * - rem_k h_{t_k}  (cif_v1, cif_predictor.py:696-735); cu_tok = prefix sum of token_num. */
int fb200_cif_integrate(const float* hidden, int ld_h, const float* alphas, const int32_t* cu_rows,
                        const int32_t* cu_tok, int batch, int total_tok, int dim, float tail_threshold,
                        const int32_t* fire_pos, const float* fire_rem, float* embeds, int ld_e,
                        void* stream);

/* ------------------------------------------------------------------ time-stamp head ---- */
/* BiCifParaformer (funasr/models/bicif_paraformer/model.py:135-142, :307-309): CifPredictorV3.get_upsample_timestamp
 * (funasr/models/bicif_paraformer/cif_predictor.py:248-293) and the fire search of ts_prediction_lfr6_standard
 * (funasr/utils/timestamp_tools.py:31-53).  The upsampling ConvTranspose1d and the LSTM input projections are
 * fb200_gemm / fb200_gemm_ex launches; these two entries are the recurrent part and the fires. */

/* Bidirectional one-layer LSTM (torch.nn.LSTM gate order i, f, g, o; zero initial state) over a packed batch, both
 * directions in one cooperative launch, fused with the Linear(2H -> 1) that follows it:
 *   xg    f32 [rows3, ld_xg]   W_ih x + b_ih + b_hh, columns [direction][gate][H] (direction 1 = reverse)
 *   utterance b owns rows [upsample*cu_rows[b], upsample*cu_rows[b+1]); each direction starts at its own end
 *   w_hh  f32 [2][4H][H], w_out f32 [2][H] (the Linear's weight split by direction)
 *   part  f32 [fb200_lstm_parts][rows3]  partial dot products: the Linear's output is bias + the sum over parts
 *   out_h f32 [rows3, 2H] or NULL        the LSTM output itself (tests)
 *   workspace: fb200_lstm_workspace_bytes(batch, H) bytes of device memory (state double buffer, cell state, barrier).
 * impl 0: the recurrent product on the tensor cores (mma.sync, bf16 hi + mid operands, 3 terms, fp32 accumulate; the state
 * travels between CTAs as bf16 hi + mid); impl 1: fp32 on the CUDA cores (check kernel).  H = 512 only; batch <= 1024. */
int fb200_lstm_workspace_bytes(int batch, int hidden, long long* bytes);
int fb200_lstm_parts(int hidden, int* parts);
int fb200_lstm_bidir(const float* xg, int ld_xg, const int32_t* cu_rows, int upsample, int batch, int rows3, int max_len3,
                     int hidden, const float* w_hh, const float* w_out, float* workspace, long long workspace_bytes,
                     float* part, float* out_h, int impl, void* stream);

/* fb200_lstm_bidir for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, kernels and results as fb200_lstm_bidir, minus the workspace: the scratch (state double buffer, cell
 * state, step counters) is a file-scope device array inside the library, sized for fb200_lstm_static_max_batch() = 64
 * utterances (786 688 bytes, held by every GPU that loads the library until the process exits).  The call allocates
 * nothing -- no cudaMalloc / cudaFree, no allocating library routine: it takes the array's device address once per
 * device (cudaGetSymbolAddress, with that device current), clears the part it needs with cudaMemsetAsync on `stream`
 * and launches.  batch > 64 is rejected with FB200_EINVAL (use fb200_lstm_bidir).  rows3 == 0 performs the per-device
 * set-up only (call it once ahead of a graph capture).  Every call shares the one array, which still holds the previous
 * call's contents: calls that overlap on the GPU (different streams without a dependency, or two threads) RACE --
 * serialise them on one stream or with events. */
int fb200_lstm_static_max_batch(void);
int fb200_lstm_bidir_static(const float* xg, int ld_xg, const int32_t* cu_rows, int upsample, int batch, int rows3,
                            int max_len3, int hidden, const float* w_hh, const float* w_out, float* part, float* out_h,
                            int impl, void* stream);

/* alphas2 = relu(sigmoid(bias + sum(part)) * smooth - noise), scaled so each utterance sums to token_num[b]
 * (cif_predictor.py:270-283) -> us_alphas [rows3]; then the sequential fp32 integrate-and-fire of cif_wo_hidden with
 * `threshold` (:71-93) -> us_peaks (optional) and the fired frame indices, relative to the utterance, in fire_idx from
 * upsample*cu_rows[b]; fire_count[b] of them.  num_chars[b] = ids of the utterance not in drop_ids (3 HOST ints);
 * when fire_count != num_chars + 1 the alphas are divided by sum / (num_chars + 1) and fired again
 * (timestamp_tools.py:46-53). */
int fb200_us_fires(const float* part, int nparts, int rows3, float bias, float smooth, float noise, float threshold,
                   const int32_t* cu_rows, int upsample, int batch, const int32_t* token_num, const int32_t* ids,
                   const int32_t* cu_tok, const int32_t* drop_ids, float* us_alphas, float* us_peaks, int32_t* fire_idx,
                   int32_t* fire_count, int32_t* num_chars, void* stream);

/* SeacoParaformer `_merge_res` on the greedy picks (funasr/models/seaco_paraformer/model.py:312-322, seaco_weight = 1):
 * out[i] = hw_ids[i] == no_bias ? dec : hw, for the ids and (when out_logp != NULL) their log-probabilities. */
int fb200_merge_hotword_ids(const int32_t* dec_ids, const float* dec_logp, const int32_t* hw_ids, const float* hw_logp,
                            int no_bias, int n, int32_t* out_ids, float* out_logp, void* stream);

/* SeacoParaformer attention-score filter (funasr/models/seaco_paraformer/model.py:268-274; `get_attn_mat`,
 * funasr/models/paraformer/decoder.py:123-137): scores[m] = sum over heads and the utterance's tokens of
 * softmax_m(scale * q_token,head . k_m,head) for the hot-word rows m of utterance b ([cu_mem[b], cu_mem[b+1]), at most 2048).
 * q f32 [tokens, ld_q], k f32 [rows, ld_k] (projected queries / keys, heads side by side); fixed summation order. */
int fb200_asf_scores(const float* q, int ld_q, const float* k, int ld_k, const int32_t* cu_tok, const int32_t* cu_mem, int batch,
                     int max_words, int heads, int head_dim, float scale, float* scores, void* stream);

/* ------------------------------------------------------------------ decoding ----------- */

/* Per row: id = argmax(logits), score = max(log_softmax) -- the greedy branch of
 * Paraformer.inference (funasr/models/paraformer/model.py:541-546) / CTC.argmax
 * (funasr/models/ctc/ctc.py:186-197). */
int fb200_argmax_logprob(const float* logits, int ld, int rows, int vocab, int32_t* ids, float* score,
                         void* stream);

/* The vocabulary / CTC projection FUSED with that greedy pick: ids[r] = argmax_c (a . w^T + bias)[r, c] (first index on
 * ties), score[r] = max log_softmax, without the [rows, vocab] logits matrix ever being written (SenseVoiceSmall: 410 MB
 * per 4096 rows).  The tcgen05 GEMM's epilogue keeps, per row and column tile, a running (max, first column, sum of
 * exponentials); a second small kernel folds the tiles.  Replaces `decoder.output_layer` + `log_softmax` + `argmax`
 * (funasr/models/paraformer/model.py:527-546) and `ctc_lo` + `CTC.argmax` (funasr/models/ctc/ctc.py:173-197).
 *   a, w      bf16 split operands as for fb200_gemm (1 or 2 parts); bias f32 [n] or NULL (-inf entries ban an id)
 *   workspace device memory of fb200_gemm_argmax_workspace(m, n) bytes, 16-byte aligned
 *   score     f32 [m] or NULL */
int fb200_gemm_argmax_workspace(int m, int n, long long* bytes);
int fb200_gemm_argmax(const void* a, int a_rows, int a_parts, const void* w, int w_parts, int m, int n, int k,
                      const float* bias, void* workspace, long long workspace_bytes, int32_t* ids, float* score,
                      void* stream);

/* CTC greedy collapse per utterance: unique_consecutive then drop ids < n_drop... drop set is
 * {blank} (SenseVoiceSmall, funasr/models/sense_voice/model.py:895-906) or {0,1,2} given as a
 * bit mask over ids 0..31 (funasr/models/ctc/model.py:243-247).  out_ids packed at cu_rows[b],
 * out_len i32 [batch]. */
int fb200_ctc_collapse(const int32_t* ids, const int32_t* cu_rows, int batch, uint32_t drop_mask,
                       int collapse_repeats, int32_t* out_ids, int32_t* out_len, void* stream);

/* ------------------------------------------------------------------ request stack ------ */

/* SMALL requests (one or a few utterances: BASELINE configs[0]) as ONE persistent cooperative kernel instead of ~580
 * launches: between fb200_stack_begin and fb200_stack_end the calls fb200_gemm (impl 0, <= 2 parts, <= 3 row taps),
 * fb200_layernorm, fb200_fsmn and fb200_attention made by THIS thread launch nothing -- they are appended to a program
 * (operand pointers, tensor maps and shapes baked in; any other fb200_* call still launches and is not part of it).
 * fb200_stack_run launches the program: one CTA per SM, tensor memory / barriers / shared-memory rings set up once, the
 * ops walked in order with a device-wide barrier wherever an op reads or overwrites what an op since the previous
 * barrier wrote or read (decided from the byte ranges at record time; the FSMN and attention of a layer share a phase).
 * Unsplit GEMM, LayerNorm and FSMN results are bit-identical to the per-op kernels'; a product consumed only by the
 * LayerNorm call right behind it is cut into K ranges which that LayerNorm sums in a fixed order; attention runs the same
 * hi.mid + mid.hi + hi.hi bf16 products on mma.sync over (utterance, head, <= 16 query rows) items (last-bit differences
 * to the tcgen05 kernel in both cases; tests/test_gpu_stack.py, ids of configs[0] == the reference's golden ids).
 * Replaces, for small requests, the per-module walk of SANMEncoder.forward (funasr/models/sanm/encoder.py:361-430) and
 * ParaformerSANMDecoder.forward (funasr/models/paraformer/decoder.py:395-404): the reference runs those as ~1700 eager
 * torch kernels per request.
 *   dim_values[ndims]  row / batch CAPACITIES used while recording (pairwise distinct, ndims <= 8): every row or batch
 *                      count of a recorded call that equals dim_values[s] becomes run-time dimension s;
 *                      fb200_stack_run(dims) then supplies the request's actual values (0 <= dims[s] <= dim_values[s]).
 *                      Buffers must be sized for the capacities; rows past dims[s] hold garbage after a run.
 * A program may not run concurrently with itself; the recorded buffers must outlive it. */
int fb200_stack_begin(void** handle, const int32_t* dim_values, int ndims);
int fb200_stack_end(void* handle);
int fb200_stack_info(void* handle, int32_t* n_ops, int32_t* n_barriers);
int fb200_stack_run(void* handle, const int32_t* dims, int ndims, void* stream);
int fb200_stack_destroy(void* handle);
/* Per-op timeline of a program: after fb200_stack_profile(handle, 1) every run leaves CTA 0's clock at eight points of
 * each op (0 start, 1 TMA producer done, 2 first operand stage landed, 3 last MMA committed, 4 the last CTA's body
 * done, 5 stores landed, 6 op body done, 7 barrier passed; 1..5 GEMM ops only).  fb200_stack_profile_read waits for
 * the device and returns the number of ops n, their kinds (0 GEMM, 1 LayerNorm, 2 FSMN, 3 attention), whether a
 * device-wide barrier follows, stamp_us[8 * n]: microseconds since the first op's start (-1: not taken), and the SM
 * clock the run saw (sm_mhz: CTA 0's cycle counter over the stamped wall time).
 * Measurement aid, no counterpart in the reference. */
int fb200_stack_profile(void* handle, int enable);
int fb200_stack_profile_read(void* handle, int32_t* kinds, int32_t* barrier_after, double* stamp_us, int capacity,
                             double* sm_mhz);

/* ------------------------------------------------------------------ native offline runtime ---- */

/* The packed Paraformer path orchestrated in C++ inside this library (csrc/runtime.cu) -- no Python, no torch: the
 * B200 counterpart of the reference's C++ offline runtime API (runtime/onnxruntime/include/funasrruntime.h:99-114).
 *   reference                                   here
 *   FunOfflineInit(model_path, threads, gpu)     fb200_offline_init(model_dir, device, precision, max_rows)
 *   FunOfflineInferBuffer(h, pcm, n_bytes, ...)  fb200_offline_infer_buffer(h, pcm16, n_bytes, 16000)   one recording
 *   (batched: paraformer-torch.cpp:301-470)      fb200_offline_infer(h, waves, lens, count)             a packed batch
 *   FunASRGetRetNumber / FunASRGetResult         fb200_result_count / fb200_result_text (+ fb200_result_ids)
 *   FunASRGetStamp                               fb200_result_stamps
 *   FunASRFreeResult / FunOfflineUninit          fb200_free_result / fb200_offline_uninit
 * model_dir holds model.fb200 (tools/export_model.py, reference state_dict names), optionally am.mvn (Kaldi-nnet CMVN,
 * parsed as paraformer.cpp:314-349) and tokens.json (JSON array of token strings).  precision: 0 = fp32 parity
 * (hi + mid bf16 operands), 1 = bf16.  waves are HOST float32 in [-1, 1] (16 kHz); every utterance is computed with
 * batch-1 semantics.  max_rows = packed encoder rows per batch (<= 0: 56832 = three waves of 148 x 128-row tiles; the
 * activation arena takes ~180 KB per row).  NULL is returned on failure (text in fb200_last_error()).  Handles are not
 * thread-safe: one handle per thread / GPU. */
void* fb200_offline_init(const char* model_dir, int device, int precision, int max_rows);
void* fb200_offline_infer(void* handle, const float* const* h_waves, const int32_t* h_lens, int count);
void* fb200_offline_infer_buffer(void* handle, const char* h_pcm16, int n_bytes, int sampling_rate);
int fb200_result_count(void* result);
const char* fb200_result_text(void* result, int index);
const int32_t* fb200_result_ids(void* result, int index, int* count);
/* Time stamps of utterance `index` (FunASRGetStamp's content, funasrruntime.h): `*count` pairs [begin ms, end ms], one per
 * id of fb200_result_ids, when the model directory holds a CifPredictorV3 head (predictor.blstm.* -- BiCifParaformer /
 * paraformer-zh); count 0 otherwise. */
const int32_t* fb200_result_stamps(void* result, int index, int* count);
void fb200_free_result(void* result);
void fb200_offline_uninit(void* handle);

#ifdef __cplusplus
}
#endif
#endif /* FUNASR_B200_H_ */
