// TORCH_LIBRARY(funasr_b200, ...) -- the thin torch custom-op layer over the C-ABI (the reference's own
// pattern for a native op is funasr/models/rwkv_bat/cuda_encoder/wkv_op.cpp:27-37).  Every op unwraps at::Tensor into raw
// device pointers, sizes and row pitches, takes the CURRENT CUDA stream of the tensor's device, calls the matching
// fb200_* entry point of libfunasr_b200.so and turns a non-zero return code into a c10::Error (Python RuntimeError) with
// fb200_last_error().  Nothing is computed here.  Outputs are caller-allocated tensors (mutable arguments), like the C-ABI's.
//
//   torch.ops.load_library("funasr_b200/libfunasr_b200_torch.so")
//   torch.ops.funasr_b200.layernorm(x, gamma, beta, 1e-12, None, xs, 2, 0, None)
//
// Built in-tree by funasr_b200/build.py (g++, torch headers; no nvcc).  funasr_b200/torch_ops.py loads it.
#include <ATen/ATen.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/library.h>

#include "../../include/funasr_b200.h"

namespace {

using at::Tensor;
using OptTensor = c10::optional<Tensor>;

// CPU tensors fail loudly BEFORE a device guard is built (the guard's own message says nothing about this library)
#define FB_GUARD(t, op)                                                                                                       \
  TORCH_CHECK((t).is_cuda(), "funasr_b200::" op ": tensor on ", (t).device(), " -- this path runs on CUDA only, no CPU fallback"); \
  c10::cuda::CUDAGuard guard((t).device())

void* stream_of(const Tensor& t) { return (void*)c10::cuda::getCurrentCUDAStream(t.get_device()).stream(); }
void ok(int rc, const char* op) { TORCH_CHECK(rc == 0, "funasr_b200::", op, " failed (", rc, "): ", fb200_last_error()); }
void cuda(const Tensor& t, at::ScalarType dt, const char* op) {
  TORCH_CHECK(t.is_cuda(), "funasr_b200::", op, ": a tensor is not on a CUDA device (this path has no CPU fallback)");
  TORCH_CHECK(t.scalar_type() == dt, "funasr_b200::", op, ": dtype ", t.scalar_type(), ", expected ", dt);
  TORCH_CHECK(t.dim() == 0 || t.stride(-1) == 1 || t.size(-1) <= 1, "funasr_b200::", op, ": rows must be contiguous");
}
const float* f32(const OptTensor& t, const char* op) {
  if (!t.has_value()) return nullptr;
  cuda(*t, at::kFloat, op);
  return t->data_ptr<float>();
}
float* f32w(const OptTensor& t, const char* op) { return const_cast<float*>(f32(t, op)); }
const int32_t* i32(const OptTensor& t, const char* op) {
  if (!t.has_value()) return nullptr;
  cuda(*t, at::kInt, op);
  return t->data_ptr<int32_t>();
}
void* bf16(const OptTensor& t, const char* op) {
  if (!t.has_value()) return nullptr;
  cuda(*t, at::kBFloat16, op);
  return t->data_ptr();
}
int pitch(const OptTensor& t) { return t.has_value() && t->dim() >= 2 ? (int)t->stride(0) : 0; }

// ---- frontend
void fbank(const Tensor& wave, const Tensor& cu_samples, const Tensor& cu_frames, int64_t total_frames, const Tensor& window,
           int64_t frame_shift, int64_t nfft, const Tensor& mel_bank, double upscale, double preemph, Tensor out) {
  FB_GUARD(wave, "fbank");
  cuda(wave, at::kFloat, "fbank"); cuda(cu_samples, at::kLong, "fbank"); cuda(cu_frames, at::kInt, "fbank");
  cuda(window, at::kFloat, "fbank"); cuda(mel_bank, at::kFloat, "fbank"); cuda(out, at::kFloat, "fbank");
  ok(fb200_fbank(wave.data_ptr<float>(), cu_samples.data_ptr<int64_t>(), cu_frames.data_ptr<int32_t>(), (int)cu_frames.numel() - 1,
                 (int)total_frames, window.data_ptr<float>(), (int)window.numel(), (int)frame_shift, (int)nfft,
                 mel_bank.data_ptr<float>(), (int)mel_bank.size(0), (float)upscale, (float)preemph, out.data_ptr<float>(),
                 stream_of(wave)), "fbank");
}

void lfr_cmvn(const Tensor& fbank_feats, const Tensor& cu_frames, const Tensor& cu_rows, int64_t total_rows, int64_t lfr_m,
              int64_t lfr_n, const OptTensor& shift, const OptTensor& scale, const OptTensor& pe_inv, double pe_scale,
              int64_t row_skip, Tensor out) {
  FB_GUARD(out, "lfr_cmvn");
  cuda(fbank_feats, at::kFloat, "lfr_cmvn"); cuda(out, at::kFloat, "lfr_cmvn");
  ok(fb200_lfr_cmvn(fbank_feats.data_ptr<float>(), i32(cu_frames, "lfr_cmvn"), i32(cu_rows, "lfr_cmvn"), (int)cu_rows.numel() - 1,
                    (int)total_rows, (int)fbank_feats.size(1), (int)lfr_m, (int)lfr_n, f32(shift, "lfr_cmvn"), f32(scale, "lfr_cmvn"),
                    f32(pe_inv, "lfr_cmvn"), (float)pe_scale, (int)row_skip, out.data_ptr<float>(), (int)out.stride(0), stream_of(out)),
     "lfr_cmvn");
}

// ---- row-wise
void layernorm(const Tensor& x, const Tensor& gamma, const Tensor& beta, double eps, const OptTensor& out_f32,
               const OptTensor& out_split, int64_t nparts, int64_t halo, const OptTensor& cu_rows) {
  FB_GUARD(x, "layernorm");
  cuda(x, at::kFloat, "layernorm"); cuda(gamma, at::kFloat, "layernorm"); cuda(beta, at::kFloat, "layernorm");
  ok(fb200_layernorm(x.data_ptr<float>(), (int)x.stride(0), (int)x.size(0), (int)x.size(1), gamma.data_ptr<float>(),
                     beta.data_ptr<float>(), (float)eps, f32w(out_f32, "layernorm"), pitch(out_f32), bf16(out_split, "layernorm"),
                     (int)nparts, (int)halo, i32(cu_rows, "layernorm"), cu_rows.has_value() ? (int)cu_rows->numel() - 1 : 0,
                     stream_of(x)), "layernorm");
}

void split_rows(const Tensor& x, Tensor out_split, int64_t nparts) {
  FB_GUARD(x, "split_rows");
  cuda(x, at::kFloat, "split_rows");
  ok(fb200_split_rows(x.data_ptr<float>(), (int)x.stride(0), (int)x.size(0), (int)x.size(1), bf16(out_split, "split_rows"),
                      (int)nparts, stream_of(x)), "split_rows");
}

void fsmn(const Tensor& v, const Tensor& cu_rows, const Tensor& weight, const OptTensor& residual, Tensor out) {
  FB_GUARD(v, "fsmn");
  cuda(v, at::kFloat, "fsmn"); cuda(weight, at::kFloat, "fsmn"); cuda(out, at::kFloat, "fsmn");
  ok(fb200_fsmn(v.data_ptr<float>(), (int)v.stride(0), i32(cu_rows, "fsmn"), (int)cu_rows.numel() - 1, (int)v.size(0), (int)v.size(1),
                weight.data_ptr<float>(), (int)weight.size(-1), f32(residual, "fsmn"), pitch(residual), out.data_ptr<float>(),
                (int)out.stride(0), stream_of(v)), "fsmn");
}

// ---- dense contractions
void gemm(const Tensor& a, int64_t a_parts, const Tensor& w, int64_t w_parts, int64_t m, int64_t n, int64_t k, int64_t taps,
          const OptTensor& bias, int64_t relu, const OptTensor& res1, const OptTensor& out_f32, int64_t f32_col0,
          const OptTensor& out_split, int64_t out_parts, int64_t impl) {
  FB_GUARD(a, "gemm");
  ok(fb200_gemm(bf16(a, "gemm"), (int)a.size(0), (int)a_parts, bf16(w, "gemm"), (int)w_parts, (int)m, (int)n, (int)k, (int)taps,
                f32(bias, "gemm"), (int)relu, f32(res1, "gemm"), pitch(res1), f32w(out_f32, "gemm"), pitch(out_f32), (int)f32_col0,
                bf16(out_split, "gemm"), (int)out_parts, (int)impl, stream_of(a)), "gemm");
}

void attention_plan(const Tensor& cu_q, const Tensor& cu_kv, Tensor plan) {
  FB_GUARD(plan, "attention_plan");
  cuda(plan, at::kInt, "attention_plan");
  ok(fb200_attention_plan(i32(cu_q, "attention_plan"), i32(cu_kv, "attention_plan"), (int)cu_q.numel() - 1, plan.data_ptr<int32_t>(),
                          (int)(plan.numel() - 4) / 4, stream_of(plan)), "attention_plan");
}

void attention(const Tensor& q, int64_t q_col0, int64_t q_part_stride, const Tensor& k, int64_t k_col0, const Tensor& v,
               int64_t v_col0, int64_t kv_part_stride, int64_t nparts, const Tensor& cu_q, const Tensor& cu_kv, const OptTensor& plan,
               int64_t max_q, int64_t heads, int64_t head_dim, double scale, const OptTensor& out_f32, const OptTensor& out_split,
               int64_t out_parts, int64_t impl) {
  FB_GUARD(q, "attention");
  ok(fb200_attention(bf16(q, "attention"), (int)q.stride(0), (int)q_col0, (int)q_part_stride, bf16(k, "attention"), (int)k.stride(0),
                     (int)k_col0, bf16(v, "attention"), (int)v.stride(0), (int)v_col0, (int)kv_part_stride, (int)nparts,
                     i32(cu_q, "attention"), i32(cu_kv, "attention"), i32(plan, "attention"), (int)cu_q.numel() - 1, (int)q.size(0),
                     (int)k.size(0), (int)max_q, (int)heads, (int)head_dim, (float)scale, f32w(out_f32, "attention"), pitch(out_f32),
                     bf16(out_split, "attention"), (int)out_parts, (int)impl, stream_of(q)), "attention");
}

// ---- CIF predictor
void cif_alpha(const Tensor& conv_relu, const Tensor& cu_rows, int64_t total_rows, const Tensor& w, const Tensor& b, double smooth,
               double noise, Tensor alphas) {
  FB_GUARD(conv_relu, "cif_alpha");
  cuda(conv_relu, at::kFloat, "cif_alpha"); cuda(alphas, at::kFloat, "cif_alpha");
  ok(fb200_cif_alpha(conv_relu.data_ptr<float>(), (int)conv_relu.stride(0), i32(cu_rows, "cif_alpha"), (int)cu_rows.numel() - 1,
                     (int)total_rows, (int)conv_relu.size(1), f32(w, "cif_alpha"), f32(b, "cif_alpha"), (float)smooth, (float)noise,
                     alphas.data_ptr<float>(), stream_of(alphas)), "cif_alpha");
}

void cif_scan(const Tensor& alphas, const Tensor& cu_rows, double tail_threshold, Tensor token_num, Tensor fire_pos, Tensor fire_rem,
              const OptTensor& fires) {
  FB_GUARD(alphas, "cif_scan");
  cuda(alphas, at::kFloat, "cif_scan"); cuda(token_num, at::kInt, "cif_scan"); cuda(fire_pos, at::kInt, "cif_scan");
  cuda(fire_rem, at::kFloat, "cif_scan");
  ok(fb200_cif_scan(alphas.data_ptr<float>(), i32(cu_rows, "cif_scan"), (int)cu_rows.numel() - 1, (float)tail_threshold,
                    token_num.data_ptr<int32_t>(), fire_pos.data_ptr<int32_t>(), fire_rem.data_ptr<float>(), f32w(fires, "cif_scan"),
                    stream_of(alphas)), "cif_scan");
}

void cif_integrate(const Tensor& hidden, const Tensor& alphas, const Tensor& cu_rows, const Tensor& cu_tok, double tail_threshold,
                   const Tensor& fire_pos, const Tensor& fire_rem, Tensor embeds) {
  FB_GUARD(hidden, "cif_integrate");
  cuda(hidden, at::kFloat, "cif_integrate"); cuda(embeds, at::kFloat, "cif_integrate");
  ok(fb200_cif_integrate(hidden.data_ptr<float>(), (int)hidden.stride(0), f32(alphas, "cif_integrate"), i32(cu_rows, "cif_integrate"),
                         i32(cu_tok, "cif_integrate"), (int)cu_rows.numel() - 1, (int)embeds.size(0), (int)hidden.size(1),
                         (float)tail_threshold, i32(fire_pos, "cif_integrate"), f32(fire_rem, "cif_integrate"),
                         embeds.data_ptr<float>(), (int)embeds.stride(0), stream_of(hidden)), "cif_integrate");
}

// ---- decoding
void argmax_logprob(const Tensor& logits, int64_t vocab, Tensor ids, Tensor score) {
  FB_GUARD(logits, "argmax_logprob");
  cuda(logits, at::kFloat, "argmax_logprob"); cuda(ids, at::kInt, "argmax_logprob"); cuda(score, at::kFloat, "argmax_logprob");
  ok(fb200_argmax_logprob(logits.data_ptr<float>(), (int)logits.stride(0), (int)logits.size(0), (int)vocab, ids.data_ptr<int32_t>(),
                          score.data_ptr<float>(), stream_of(logits)), "argmax_logprob");
}

void ctc_collapse(const Tensor& ids, const Tensor& cu_rows, int64_t drop_mask, int64_t collapse_repeats, Tensor out_ids, Tensor out_len) {
  FB_GUARD(ids, "ctc_collapse");
  cuda(out_ids, at::kInt, "ctc_collapse"); cuda(out_len, at::kInt, "ctc_collapse");
  ok(fb200_ctc_collapse(i32(ids, "ctc_collapse"), i32(cu_rows, "ctc_collapse"), (int)cu_rows.numel() - 1, (uint32_t)drop_mask,
                        (int)collapse_repeats, out_ids.data_ptr<int32_t>(), out_len.data_ptr<int32_t>(), stream_of(ids)),
     "ctc_collapse");
}

int64_t version() { return fb200_version(); }

}  // namespace

TORCH_LIBRARY(funasr_b200, m) {
  m.def("version() -> int", &version);
  m.def("fbank(Tensor wave, Tensor cu_samples, Tensor cu_frames, int total_frames, Tensor window, int frame_shift, int nfft, "
        "Tensor mel_bank, float upscale, float preemph, Tensor(a!) out) -> ()", &fbank);
  m.def("lfr_cmvn(Tensor fbank, Tensor cu_frames, Tensor cu_rows, int total_rows, int lfr_m, int lfr_n, Tensor? shift, Tensor? scale, "
        "Tensor? pe_inv, float pe_scale, int row_skip, Tensor(a!) out) -> ()", &lfr_cmvn);
  m.def("layernorm(Tensor x, Tensor gamma, Tensor beta, float eps, Tensor(a!)? out_f32, Tensor(b!)? out_split, int nparts, int halo, "
        "Tensor? cu_rows) -> ()", &layernorm);
  m.def("split_rows(Tensor x, Tensor(a!) out_split, int nparts) -> ()", &split_rows);
  m.def("fsmn(Tensor v, Tensor cu_rows, Tensor weight, Tensor? residual, Tensor(a!) out) -> ()", &fsmn);
  m.def("gemm(Tensor a, int a_parts, Tensor w, int w_parts, int m, int n, int k, int taps, Tensor? bias, int relu, Tensor? res1, "
        "Tensor(a!)? out_f32, int f32_col0, Tensor(b!)? out_split, int out_parts, int impl) -> ()", &gemm);
  m.def("attention_plan(Tensor cu_q, Tensor cu_kv, Tensor(a!) plan) -> ()", &attention_plan);
  m.def("attention(Tensor q, int q_col0, int q_part_stride, Tensor k, int k_col0, Tensor v, int v_col0, int kv_part_stride, int nparts, "
        "Tensor cu_q, Tensor cu_kv, Tensor? plan, int max_q, int heads, int head_dim, float scale, Tensor(a!)? out_f32, "
        "Tensor(b!)? out_split, int out_parts, int impl) -> ()", &attention);
  m.def("cif_alpha(Tensor conv_relu, Tensor cu_rows, int total_rows, Tensor w, Tensor b, float smooth, float noise, Tensor(a!) alphas) -> ()",
        &cif_alpha);
  m.def("cif_scan(Tensor alphas, Tensor cu_rows, float tail_threshold, Tensor(a!) token_num, Tensor(b!) fire_pos, Tensor(c!) fire_rem, "
        "Tensor(d!)? fires) -> ()", &cif_scan);
  m.def("cif_integrate(Tensor hidden, Tensor alphas, Tensor cu_rows, Tensor cu_tok, float tail_threshold, Tensor fire_pos, "
        "Tensor fire_rem, Tensor(a!) embeds) -> ()", &cif_integrate);
  m.def("argmax_logprob(Tensor logits, int vocab, Tensor(a!) ids, Tensor(b!) score) -> ()", &argmax_logprob);
  m.def("ctc_collapse(Tensor ids, Tensor cu_rows, int drop_mask, int collapse_repeats, Tensor(a!) out_ids, Tensor(b!) out_len) -> ()",
        &ctc_collapse);
}
