// Request-stack kernel (stack_sm100.cu): the per-op entry points (fb200_gemm, fb200_layernorm, fb200_fsmn,
// fb200_attention) append to a PROGRAM instead of launching while a recording is open on the calling thread; the program
// later runs as ONE persistent cooperative kernel (grid barriers between dependent ops).  This header is the seam between
// the op translation units and the recorder.
#pragma once
#include "common.cuh"

namespace fb200 {

constexpr int kMaxTerms = 6;
constexpr int kMaxTaps = 24;

struct GemmParams {
  int m, n, k_blocks, kp_a, kp_w;
  int taps, nterms;
  int term_a[kMaxTerms], term_b[kMaxTerms];
  int a_col0;                                // first A column; tap j reads A row m + tap_row[j], columns + tap_col[j]
  int tap_row[kMaxTaps], tap_col[kMaxTaps];
  // Column groups (0 = none): output columns [g * group_n, (g + 1) * group_n) are an independent product of the A
  // columns shifted by g * a_col_step with the weight rows starting at g * w_group_rows -- the per-head (q+v).p^T
  // products of the rel-pos attention as ONE launch.
  int group_n, a_col_step, w_group_rows;
  int tiles_m, tiles_n;
  const float* bias;
  int act;  // 0 none, 1 ReLU, 2 swish y * sigmoid(y)
  const float* res1;
  int ld_res1;
  float* out_f32;
  int ld_out, f32_col0;  // out_f32 holds columns [f32_col0, n) only
  __nv_bfloat16* out_split;
  int out_parts, out_kp;
  // fused greedy pick (fb200_gemm_argmax): no matrix is written; every column tile leaves, per row, its maximum, the first
  // column that reaches it and sum(exp(y - max)) in amax_part[row * tiles_n + tile] = (max, sum, column as bits, -)
  float4* amax_part;
  int direct_store;  // epilogue writes its rows with plain 16-byte stores instead of staging boxes + TMA stores
};

// true while fb200_stack_begin() ... fb200_stack_end() is open on this thread
bool stack_recording();
int stack_record_gemm(const void* a, int a_rows, int lda, int a_parts, const void* w, int w_rows, int ldw, const GemmParams& p);
int stack_record_layernorm(const float* x, int ld_x, int rows, int dim, const float* gamma, const float* beta, float eps,
                           float* out_f32, int ld_f32, void* out_split, int nparts, int halo, const int32_t* cu_rows,
                           int batch);
int stack_record_fsmn(const float* v, int ld_v, const int32_t* cu_rows, int batch, int total_rows, int dim,
                      const float* weight, int ksize, const float* residual, int ld_res, float* out, int ld_out);
int stack_record_attention(const void* q, int ld_q, int q_col0, int q_part_stride, const void* k, int ld_k, int k_col0,
                           const void* v, int ld_v, int v_col0, int kv_part_stride, int nparts, const int32_t* cu_q,
                           const int32_t* cu_kv, int batch, int q_rows, int kv_rows, int heads, int head_dim, float scale,
                           float* out_f32, int ld_out, void* out_split, int out_parts);

}  // namespace fb200
