// Native offline recogniser: the packed Paraformer path (frontend -> SAN-M encoder -> CIF predictor -> SAN-M decoder
// -> greedy pick) orchestrated in C++ over the kernels of this library -- no Python, no torch.  It is the B200
// counterpart of the reference's C++ offline runtime (runtime/onnxruntime/include/funasrruntime.h:99-114:
// FunOfflineInit / FunOfflineInferBuffer / FunASRGetResult / FunASRGetRetNumber / FunASRFreeResult / FunOfflineUninit,
// implemented over ONNX Runtime / libtorch in runtime/onnxruntime/src/paraformer.cpp and paraformer-torch.cpp).
//
// Model directory (data formats either side of the path):
//   model.fb200   weights under the reference's state_dict names, fp32, written by tools/export_model.py from a
//                 model.pt / state_dict (the reference's C++ runtime reads model.pt through libtorch, which this
//                 library does not link); also carries the frontend tables (window, mel bank) the exporter computed
//   am.mvn        Kaldi-nnet CMVN text, parsed here like runtime/onnxruntime/src/paraformer.cpp:314-349
//   tokens.json   optional JSON array of token strings (runtime/onnxruntime/src/vocab.cpp); without it the result
//                 text is the space-separated token ids
//
// The op sequence is the one of funasr_b200/engine.py (ParaformerEngine.forward_packed), which cites the reference
// model code line by line; tests/test_runtime.py checks that both produce identical token ids.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <fstream>
#include <map>
#include <numeric>
#include <sstream>
#include <string>
#include <thread>
#include <tuple>
#include <vector>

#include "common.cuh"

namespace fb200 {
namespace rt {

typedef __nv_bfloat16 bf16;

#define RT_CUDA(call)                                                                   \
  do {                                                                                  \
    cudaError_t e_ = (call);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      set_error("runtime: %s failed: %s", #call, cudaGetErrorString(e_));               \
      return FB200_ECUDA;                                                               \
    }                                                                                   \
  } while (0)
#define RT_OK(call)            \
  do {                         \
    int rc_ = (call);          \
    if (rc_ != FB200_OK) return rc_; \
  } while (0)

static inline int kp64(int k) { return (k + 63) / 64 * 64; }

// Bump allocator over one device slab: per-batch activations (reset before every batch).
struct Arena {
  char* base = nullptr;
  size_t cap = 0, off = 0;
  template <typename T>
  T* get(size_t n) {  // never fails; the caller checks ok() before touching what it was given
    size_t bytes = (n * sizeof(T) + 255) / 256 * 256;
    T* p = reinterpret_cast<T*>(base + off);
    off += bytes;
    return p;
  }
  bool ok() const { return off <= cap; }
};

struct HostTensor {
  std::vector<int64_t> shape;
  std::vector<float> data;
  int64_t numel() const { return std::accumulate(shape.begin(), shape.end(), (int64_t)1, std::multiplies<int64_t>()); }
};

struct Linear {  // bf16 split weight [taps * n, nparts * Kp] + fp32 bias
  bf16* w = nullptr;
  float* bias = nullptr;
  int n = 0, k = 0, taps = 1;
};
struct Norm {
  float *g = nullptr, *b = nullptr;
};
struct EncLayer {
  Norm norm1, norm2;
  Linear qkv, out, w1, w2;
  float* fsmn = nullptr;
  int fsmn_k = 0, din = 0;
};
struct DecLayer {
  Norm norm1, ffn_norm, norm2, norm3;
  Linear w1, w2, q, out;
  float* fsmn = nullptr;
  int fsmn_k = 0;
  bool has_fsmn = false, has_src = false;
  int src_index = -1;
};

struct Result {
  std::vector<std::vector<int32_t>> ids;
  std::vector<std::string> text;
  std::vector<std::vector<int32_t>> stamps;  // [begin ms, end ms] per kept token (models with the CifPredictorV3 head)
};

// fired upsampled frames -> [begin ms, end ms] per token: the arithmetic of ts_prediction_lfr6_standard
// (funasr/utils/timestamp_tools.py:31-106) in double, as the reference (numpy int64 + Python floats) computes it
static void stamps_from_fires(const int32_t* fire_idx, int n, int num_frames, int num_chars, std::vector<int32_t>* out) {
  out->clear();
  if (num_chars <= 0 || n < 2) return;
  const double kShift = -1.5, kRate = 10.0 * 6 / 1000 / 3;
  const int kStartEnd = 5, kMaxDur = 12;
  const int m = std::min(n - 1, num_chars);
  for (int i = 0; i < m; ++i) {
    const double f0 = (double)fire_idx[i] + kShift, f1 = (double)fire_idx[i + 1] + kShift;
    const bool split = f1 - f0 > kMaxDur;
    const double begin = f0 * kRate;
    double end = split ? (f0 + kMaxDur) * kRate : f1 * kRate;
    if (i == n - 2 && !split) {  // the last list entry is the last token (otherwise the silence after it)
      if (num_frames - f1 > kStartEnd) end = ((num_frames + f1) * 0.5) * kRate;
      else end = num_frames * kRate;
    }
    out->push_back((int32_t)(begin * 1000));
    out->push_back((int32_t)(end * 1000));
  }
}

class Recognizer {
 public:
  ~Recognizer() {
    if (enc_prog_) fb200_stack_destroy(enc_prog_);
    if (dec_prog_) fb200_stack_destroy(dec_prog_);
    for (auto& g : graphs_) cudaGraphExecDestroy(g.second);
    for (void* p : owned_) cudaFree(p);
    if (arena_.base) cudaFree(arena_.base);
    for (int i = 0; i < 2; ++i) {
      if (dwave_[i]) cudaFree(dwave_[i]);
      if (copied_[i]) cudaEventDestroy(copied_[i]);
    }
    if (copy_stream_) cudaStreamDestroy(copy_stream_);
    if (stream_) cudaStreamDestroy(stream_);
  }
  int init(const std::string& dir, int device, int nparts, int max_rows);
  int recognize(const float* const* waves, const int* lens, int n, Result* res);

 private:
  int load_file(const std::string& path);
  int load_cmvn(const std::string& path);
  void load_tokens(const std::string& path);
  const HostTensor* find(const std::string& name) const {
    auto it = host_.find(name);
    return it == host_.end() ? nullptr : &it->second;
  }
  bool has(const std::string& name) const { return host_.count(name) != 0; }
  int upload(const std::vector<float>& v, float** out);
  int make_norm(const std::string& prefix, Norm* n);
  int make_linear(const std::string& wname, const std::string& bname, Linear* l);
  int make_linear_from(const std::vector<float>& w, int rows, int k, int taps, const std::vector<float>* bias, Linear* l);
  int forward(const std::vector<int>& batch, int slot, const int* lens, Result* res, const std::vector<int>* next,
              const float* const* waves);
  int stage(const std::vector<int>& batch, int slot, const float* const* waves, const int* lens);
  int reserve_wave(size_t samples);

  int device_ = 0, np_ = 2, max_rows_ = 56832, heads_ = 4;
  int d_ = 512, din_ = 560, ff_ = 2048, vocab_ = 0, n_mels_ = 80, lfr_m_ = 7, lfr_n_ = 6;
  int frame_len_ = 400, frame_shift_ = 160, nfft_ = 512;
  float eps_ = 1e-12f, tail_ = 0.45f;
  cudaStream_t stream_ = nullptr;
  std::map<std::string, HostTensor> host_;
  std::vector<void*> owned_;
  Arena arena_;
  // waveforms: caller memory -> device, double-buffered on a copy stream so that the upload of batch k + 1 runs
  // under the kernels of batch k.  The copies are issued straight from the caller's (possibly pageable) buffers: the
  // driver's own staging moved them at ~13 GB/s here, a memcpy into a pinned slab on this thread at a fraction of that
  // (measured: 498 ms/step with direct copies on one stream, 526 ms with a pinned staging slab).
  cudaStream_t copy_stream_ = nullptr;
  float* dwave_[2] = {nullptr, nullptr};
  cudaEvent_t copied_[2] = {nullptr, nullptr};
  size_t wave_cap_ = 0;
  float *window_ = nullptr, *mel_ = nullptr, *pe_inv_ = nullptr, *shift_ = nullptr, *scale_ = nullptr;
  std::vector<EncLayer> enc_;
  Norm enc_after_;
  Linear cif_conv_;
  float *cif_w_ = nullptr, *cif_b_ = nullptr;
  std::vector<DecLayer> dec_;
  Norm dec_after_;
  Linear dec_out_, dec_kv_;
  std::vector<std::string> tokens_;
  // CUDA graphs of the launch-bound stretch (frontend + encoder + predictor: ~470 launches) for shapes that come back --
  // a request of one short utterance is bound by the host's launch rate (13 us per launch, 5.6 of 7.7 ms for 10 s of audio),
  // not by the GPU.  Key = everything the launches' host-side arguments depend on; the per-utterance prefix arrays and the
  // waveforms live at fixed device addresses for a key and are refreshed outside the graph.  A key is captured the second
  // time it is seen (the first, eager pass also performs the one-time cudaFuncSetAttribute calls).
  typedef std::tuple<int, int, int, int, int, int, int> GraphKey;  // slot, utterances, frames, rows, longest utterance, tokens, longest token run
  // (every arena offset behind the fbank buffer depends on the frame count F through the 256-byte rounding, so BOTH stretches key on F)
  std::map<GraphKey, cudaGraphExec_t> graphs_;
  std::map<GraphKey, int> graph_seen_;
  bool use_graphs_ = true;
  template <typename F>
  int run_graphed(const GraphKey& key, bool allowed, F&& issue);
  // Small requests (<= kStackBatch utterances, <= kStackRows encoder rows -- one 30 s utterance; BASELINE configs[0]): the encoder and decoder
  // stretches run as two persistent cooperative kernels (fb200_stack_*, csrc/stack_sm100.cu) recorded once over
  // capacity-sized buffers of their own; FB200_STACK=0 keeps the per-op launches.
  static constexpr int kStackRows = 512, kStackTokens = 384, kStackBatch = 8;
  struct EncBufs {
    float* feats; bf16 *xs0, *xs, *h, *qkv; float *v, *mem, *x0, *x1, *enc; bf16 *halo, *encs; int32_t *d_cur, *plan;
  };
  struct DecBufs {
    bf16 *encs, *kv, *qs, *ys, *hs; float *hf, *t, *t2, *n2, *b0, *b1, *embeds, *logits; int32_t *d_cut, *d_cur, *dplan;
  };
  int encoder_ops(const EncBufs& b, int R, int B, int max_t, void* st);
  int decoder_ops(const DecBufs& b, int R, int L, int B, int max_l, int ldv, void* st);
  int small_init();
  bool use_stack_ = true, small_ready_ = false;
  EncBufs small_enc_ = {};
  DecBufs small_dec_ = {};
  void *enc_prog_ = nullptr, *dec_prog_ = nullptr;
  // CifPredictorV3 time-stamp head (BiCifParaformer), present when the checkpoint holds predictor.blstm.*
  bool has_stamps_ = false;
  static constexpr int kMaxStampBatch = 1024;
  Linear ts_up_, ts_proj_;
  float *ts_whh_ = nullptr, *ts_wout_ = nullptr, ts_bout_ = 0.f;
  int ts_parts_ = 0;
};

// ---------------------------------------------------------------- loading -----------------
int Recognizer::load_file(const std::string& path) {
  std::ifstream f(path, std::ios::binary);
  if (!f) {
    set_error("runtime: cannot open %s", path.c_str());
    return FB200_EINVAL;
  }
  f.seekg(0, std::ios::end);
  const std::streamoff file_bytes = f.tellg();
  f.seekg(0, std::ios::beg);
  char magic[8];
  uint32_t count = 0;
  f.read(magic, 8);
  f.read(reinterpret_cast<char*>(&count), 4);
  if (!f || memcmp(magic, "FB200W1\0", 8) != 0) {
    set_error("runtime: %s is not a model.fb200 file", path.c_str());
    return FB200_EINVAL;
  }
  for (uint32_t i = 0; i < count; ++i) {
    uint16_t nlen = 0;
    uint8_t ndim = 0;
    f.read(reinterpret_cast<char*>(&nlen), 2);
    std::string name(nlen, '\0');
    f.read(&name[0], nlen);
    f.read(reinterpret_cast<char*>(&ndim), 1);
    HostTensor t;
    t.shape.resize(ndim);
    f.read(reinterpret_cast<char*>(t.shape.data()), 8 * ndim);
    // a corrupt header must not turn into a huge or negative allocation: every extent non-negative, the element count
    // bounded by what is left of the file
    bool sane = (bool)f && ndim <= 8;
    unsigned long long numel = 1;
    for (size_t a = 0; sane && a < t.shape.size(); ++a) {
      sane = t.shape[a] >= 0 && (t.shape[a] == 0 || numel <= (unsigned long long)file_bytes / (unsigned long long)t.shape[a]);
      if (sane) numel *= (unsigned long long)t.shape[a];
    }
    if (sane) {
      const std::streamoff at = f.tellg();
      sane = at >= 0 && numel * 4ull <= (unsigned long long)(file_bytes - at);
    }
    if (!sane) {
      set_error("runtime: %s: bad shape at tensor %u (%s)", path.c_str(), i, name.c_str());
      return FB200_EINVAL;
    }
    t.data.resize((size_t)numel);
    f.read(reinterpret_cast<char*>(t.data.data()), 4 * (std::streamsize)t.data.size());
    if (!f) {
      set_error("runtime: %s is truncated at tensor %u (%s)", path.c_str(), i, name.c_str());
      return FB200_EINVAL;
    }
    host_[name] = std::move(t);
  }
  return FB200_OK;
}

// Kaldi-nnet am.mvn: the <LearnRateCoef> line after <AddShift> is the shift, after <Rescale> the scale; values are
// fields [3 : -1] of that line (funasr/frontends/wav_frontend.py:15-38, runtime/onnxruntime/src/paraformer.cpp:314-349).
int Recognizer::load_cmvn(const std::string& path) {
  std::ifstream f(path);
  if (!f) return FB200_EINVAL;  // optional
  std::vector<std::string> lines;
  for (std::string ln; std::getline(f, ln);) lines.push_back(ln);
  std::vector<float> shift, scale;
  for (size_t i = 0; i + 1 < lines.size(); ++i) {
    std::istringstream a(lines[i]);
    std::string tag;
    a >> tag;
    if (tag != "<AddShift>" && tag != "<Rescale>") continue;
    std::istringstream b(lines[i + 1]);
    std::vector<std::string> fld;
    for (std::string s; b >> s;) fld.push_back(s);
    if (fld.empty() || fld[0] != "<LearnRateCoef>" || fld.size() < 5) continue;
    std::vector<float>& dst = tag == "<AddShift>" ? shift : scale;
    for (size_t j = 3; j + 1 < fld.size(); ++j) dst.push_back((float)atof(fld[j].c_str()));
  }
  if (shift.empty() || shift.size() != scale.size()) {
    set_error("runtime: %s holds no <AddShift>/<Rescale> statistics", path.c_str());
    return FB200_EINVAL;
  }
  shift.resize(din_);
  scale.resize(din_);
  RT_OK(upload(shift, &shift_));
  RT_OK(upload(scale, &scale_));
  return FB200_OK;
}

// tokens.json: a JSON array of strings (\" \\ \uXXXX escapes of the BMP handled).
void Recognizer::load_tokens(const std::string& path) {
  std::ifstream f(path);
  if (!f) return;
  std::stringstream ss;
  ss << f.rdbuf();
  const std::string s = ss.str();
  size_t i = 0;
  while (i < s.size()) {
    if (s[i] != '"') {
      ++i;
      continue;
    }
    std::string tok;
    for (++i; i < s.size() && s[i] != '"'; ++i) {
      if (s[i] != '\\' || i + 1 >= s.size()) {
        tok.push_back(s[i]);
        continue;
      }
      const char c = s[++i];
      if (c == 'u' && i + 4 < s.size()) {
        const unsigned cp = (unsigned)strtoul(s.substr(i + 1, 4).c_str(), nullptr, 16);
        i += 4;
        if (cp < 0x80) tok.push_back((char)cp);
        else if (cp < 0x800) tok.push_back((char)(0xC0 | (cp >> 6))), tok.push_back((char)(0x80 | (cp & 0x3F)));
        else tok.push_back((char)(0xE0 | (cp >> 12))), tok.push_back((char)(0x80 | ((cp >> 6) & 0x3F))), tok.push_back((char)(0x80 | (cp & 0x3F)));
      } else {
        tok.push_back(c == 'n' ? '\n' : c == 't' ? '\t' : c);
      }
    }
    ++i;
    tokens_.push_back(tok);
  }
}

int Recognizer::upload(const std::vector<float>& v, float** out) {
  float* p = nullptr;
  RT_CUDA(cudaMalloc(&p, std::max<size_t>(v.size(), 1) * sizeof(float)));
  owned_.push_back(p);
  RT_CUDA(cudaMemcpyAsync(p, v.data(), v.size() * sizeof(float), cudaMemcpyHostToDevice, stream_));
  RT_CUDA(cudaStreamSynchronize(stream_));
  *out = p;
  return FB200_OK;
}

int Recognizer::make_norm(const std::string& prefix, Norm* n) {
  const HostTensor *g = find(prefix + "weight"), *b = find(prefix + "bias");
  if (!g || !b) {
    set_error("runtime: missing %sweight / bias", prefix.c_str());
    return FB200_EINVAL;
  }
  RT_OK(upload(g->data, &n->g));
  return upload(b->data, &n->b);
}

int Recognizer::make_linear_from(const std::vector<float>& w, int rows, int k, int taps, const std::vector<float>* bias, Linear* l) {
  float* tmp = nullptr;
  RT_CUDA(cudaMalloc(&tmp, w.size() * sizeof(float)));
  RT_CUDA(cudaMemcpyAsync(tmp, w.data(), w.size() * sizeof(float), cudaMemcpyHostToDevice, stream_));
  bf16* sp = nullptr;
  RT_CUDA(cudaMalloc(&sp, (size_t)rows * np_ * kp64(k) * sizeof(bf16)));
  owned_.push_back(sp);
  int rc = fb200_split_weight(tmp, rows, k, sp, np_, stream_);
  cudaStreamSynchronize(stream_);
  cudaFree(tmp);
  RT_OK(rc);
  l->w = sp;
  l->n = rows / taps;
  l->k = k;
  l->taps = taps;
  l->bias = nullptr;
  if (bias) RT_OK(upload(*bias, &l->bias));
  return FB200_OK;
}

int Recognizer::make_linear(const std::string& wname, const std::string& bname, Linear* l) {
  const HostTensor* w = find(wname);
  if (!w) {
    set_error("runtime: missing %s", wname.c_str());
    return FB200_EINVAL;
  }
  const HostTensor* b = bname.empty() ? nullptr : find(bname);
  if (w->shape.size() == 3) {  // Conv1d [out, in, taps] -> [taps][out, in]
    const int out = (int)w->shape[0], in = (int)w->shape[1], taps = (int)w->shape[2];
    std::vector<float> r((size_t)taps * out * in);
    for (int o = 0; o < out; ++o)
      for (int c = 0; c < in; ++c)
        for (int t = 0; t < taps; ++t) r[((size_t)t * out + o) * in + c] = w->data[((size_t)o * in + c) * taps + t];
    return make_linear_from(r, taps * out, in, taps, b ? &b->data : nullptr, l);
  }
  return make_linear_from(w->data, (int)w->shape[0], (int)w->shape[1], 1, b ? &b->data : nullptr, l);
}

int Recognizer::init(const std::string& dir, int device, int nparts, int max_rows) {
  device_ = device;
  np_ = nparts;
  if (max_rows < 1 || max_rows > 65536) {  // element offsets of the [rows, 2 x 16384] K / V projection stay below 2^31
    set_error("runtime: max_rows %d outside [1, 65536]", max_rows);
    return FB200_EINVAL;
  }
  max_rows_ = max_rows;
  const char* genv = getenv("FB200_RT_GRAPHS");  // FB200_RT_GRAPHS=0: always launch kernel by kernel
  use_graphs_ = genv == nullptr || atoi(genv) != 0;
  RT_OK(load_file(dir + "/model.fb200"));  // host only: a bad model directory is reported before any CUDA call
  load_tokens(dir + "/tokens.json");
  RT_CUDA(cudaSetDevice(device_));
  RT_CUDA(cudaStreamCreate(&stream_));
  RT_CUDA(cudaStreamCreate(&copy_stream_));
  for (int i = 0; i < 2; ++i) RT_CUDA(cudaEventCreateWithFlags(&copied_[i], cudaEventDisableTiming));
  const HostTensor *win = find("frontend.window"), *mel = find("frontend.mel"), *pe = find("frontend.pe_inv");
  const HostTensor* meta = find("frontend.meta");  // [lfr_m, lfr_n, frame_shift, nfft, heads, tail_threshold]
  if (!win || !mel || !pe || !meta || meta->data.size() < 6) {
    set_error("runtime: model.fb200 lacks the frontend tables (frontend.window / mel / pe_inv / meta)");
    return FB200_EINVAL;
  }
  frame_len_ = (int)win->data.size();
  n_mels_ = (int)mel->shape[0];
  lfr_m_ = (int)meta->data[0];
  lfr_n_ = (int)meta->data[1];
  frame_shift_ = (int)meta->data[2];
  nfft_ = (int)meta->data[3];
  heads_ = (int)meta->data[4];
  tail_ = meta->data[5];
  din_ = n_mels_ * lfr_m_;
  RT_OK(upload(win->data, &window_));
  RT_OK(upload(mel->data, &mel_));
  RT_OK(upload(pe->data, &pe_inv_));
  if (load_cmvn(dir + "/am.mvn") != FB200_OK) {
    const HostTensor *sh = find("frontend.cmvn_shift"), *sc = find("frontend.cmvn_scale");
    if (sh && sc) {
      RT_OK(upload(sh->data, &shift_));
      RT_OK(upload(sc->data, &scale_));
    }
  }
  // encoder
  for (const char* group : {"encoder.encoders0.", "encoder.encoders."}) {
    for (int i = 0;; ++i) {
      const std::string p = std::string(group) + std::to_string(i) + ".";
      if (!has(p + "norm1.weight")) break;
      EncLayer ly;
      RT_OK(make_norm(p + "norm1.", &ly.norm1));
      RT_OK(make_norm(p + "norm2.", &ly.norm2));
      RT_OK(make_linear(p + "self_attn.linear_q_k_v.weight", p + "self_attn.linear_q_k_v.bias", &ly.qkv));
      RT_OK(make_linear(p + "self_attn.linear_out.weight", p + "self_attn.linear_out.bias", &ly.out));
      RT_OK(make_linear(p + "feed_forward.w_1.weight", p + "feed_forward.w_1.bias", &ly.w1));
      RT_OK(make_linear(p + "feed_forward.w_2.weight", p + "feed_forward.w_2.bias", &ly.w2));
      const HostTensor* fw = find(p + "self_attn.fsmn_block.weight");
      if (!fw) {
        set_error("runtime: missing %sself_attn.fsmn_block.weight", p.c_str());
        return FB200_EINVAL;
      }
      ly.fsmn_k = (int)fw->shape.back();
      RT_OK(upload(fw->data, &ly.fsmn));
      ly.din = ly.qkv.k;
      enc_.push_back(ly);
    }
  }
  if (enc_.empty()) {
    set_error("runtime: model.fb200 holds no encoder layers");
    return FB200_EINVAL;
  }
  d_ = enc_[0].out.n;
  ff_ = enc_[0].w1.n;
  RT_OK(make_norm("encoder.after_norm.", &enc_after_));
  // predictor
  RT_OK(make_linear("predictor.cif_conv1d.weight", "predictor.cif_conv1d.bias", &cif_conv_));
  const HostTensor *cow = find("predictor.cif_output.weight"), *cob = find("predictor.cif_output.bias");
  if (!cow || !cob || (int)cow->data.size() < d_ || cob->data.empty()) {
    set_error("runtime: missing predictor.cif_output.weight / bias");
    return FB200_EINVAL;
  }
  RT_OK(upload(cow->data, &cif_w_));
  RT_OK(upload(cob->data, &cif_b_));
  if (has("predictor.blstm.weight_hh_l0") && has("predictor.upsample_cnn.weight") && has("predictor.cif_output2.weight")) {
    // bicif_paraformer/cif_predictor.py:139-145: ConvTranspose1d [in, out, k] (stride k) as a GEMM with weight rows (j, o),
    // both directions' input projections stacked, recurrent weights and the Linear(2H -> 1) kept in fp32
    const HostTensor* uw = find("predictor.upsample_cnn.weight");
    const HostTensor* ub = find("predictor.upsample_cnn.bias");
    if (uw->shape.size() != 3 || !ub) {
      set_error("runtime: predictor.upsample_cnn.weight must be [in, out, k] with a bias");
      return FB200_EINVAL;
    }
    const int cin = (int)uw->shape[0], cout = (int)uw->shape[1], k = (int)uw->shape[2];
    if (k != 3 || cout != 512 || cin != d_) {
      set_error("runtime: time-stamp head with upsample %d / hidden %d unsupported (3 / 512)", k, cout);
      return FB200_EUNSUPPORTED;
    }
    std::vector<float> w((size_t)k * cout * cin), b((size_t)k * cout);
    for (int j = 0; j < k; ++j)
      for (int o = 0; o < cout; ++o) {
        b[(size_t)j * cout + o] = ub->data[o];
        for (int c = 0; c < cin; ++c) w[((size_t)j * cout + o) * cin + c] = uw->data[((size_t)c * cout + o) * k + j];
      }
    RT_OK(make_linear_from(w, k * cout, cin, 1, &b, &ts_up_));
    std::vector<float> wih, bih, whh;
    for (const char* sfx : {"", "_reverse"}) {
      const HostTensor *wi = find(std::string("predictor.blstm.weight_ih_l0") + sfx), *wh = find(std::string("predictor.blstm.weight_hh_l0") + sfx);
      const HostTensor *bi = find(std::string("predictor.blstm.bias_ih_l0") + sfx), *bh = find(std::string("predictor.blstm.bias_hh_l0") + sfx);
      if (!wi || !wh || !bi || !bh) {
        set_error("runtime: incomplete predictor.blstm.* tensors");
        return FB200_EINVAL;
      }
      wih.insert(wih.end(), wi->data.begin(), wi->data.end());
      whh.insert(whh.end(), wh->data.begin(), wh->data.end());
      for (size_t i = 0; i < bi->data.size(); ++i) bih.push_back(bi->data[i] + bh->data[i]);
    }
    RT_OK(make_linear_from(wih, 8 * cout, cout, 1, &bih, &ts_proj_));
    RT_OK(upload(whh, &ts_whh_));
    const HostTensor *o2w = find("predictor.cif_output2.weight"), *o2b = find("predictor.cif_output2.bias");
    if (!ub || !o2w || !o2b || (int)ub->data.size() < cout || (int)o2w->data.size() < 2 * cout || o2b->data.empty()) {
      set_error("runtime: incomplete time-stamp head (predictor.upsample_cnn.bias / cif_output2.*)");
      return FB200_EINVAL;
    }
    RT_OK(upload(o2w->data, &ts_wout_));
    ts_bout_ = o2b->data[0];
    RT_OK(fb200_lstm_parts(cout, &ts_parts_));
    has_stamps_ = true;
  }
  // decoder
  std::vector<float> kvw, kvb;
  int n_src = 0;
  for (const char* group : {"decoder.decoders.", "decoder.decoders2.", "decoder.decoders3."}) {
    for (int i = 0;; ++i) {
      const std::string p = std::string(group) + std::to_string(i) + ".";
      if (!has(p + "norm1.weight")) break;
      DecLayer ly;
      RT_OK(make_norm(p + "norm1.", &ly.norm1));
      RT_OK(make_linear(p + "feed_forward.w_1.weight", p + "feed_forward.w_1.bias", &ly.w1));
      RT_OK(make_norm(p + "feed_forward.norm.", &ly.ffn_norm));
      RT_OK(make_linear(p + "feed_forward.w_2.weight", has(p + "feed_forward.w_2.bias") ? p + "feed_forward.w_2.bias" : "", &ly.w2));
      ly.has_fsmn = has(p + "self_attn.fsmn_block.weight");
      ly.has_src = has(p + "src_attn.linear_q.weight");
      if (ly.has_fsmn) {
        RT_OK(make_norm(p + "norm2.", &ly.norm2));
        const HostTensor* fw = find(p + "self_attn.fsmn_block.weight");
        ly.fsmn_k = (int)fw->shape.back();
        RT_OK(upload(fw->data, &ly.fsmn));
      }
      if (ly.has_src) {
        RT_OK(make_norm(p + "norm3.", &ly.norm3));
        RT_OK(make_linear(p + "src_attn.linear_q.weight", p + "src_attn.linear_q.bias", &ly.q));
        RT_OK(make_linear(p + "src_attn.linear_out.weight", p + "src_attn.linear_out.bias", &ly.out));
        const HostTensor *w = find(p + "src_attn.linear_k_v.weight"), *b = find(p + "src_attn.linear_k_v.bias");
        if (!w || !b) {
          set_error("runtime: missing %ssrc_attn.linear_k_v.weight / bias", p.c_str());
          return FB200_EINVAL;
        }
        kvw.insert(kvw.end(), w->data.begin(), w->data.end());
        kvb.insert(kvb.end(), b->data.begin(), b->data.end());
        ly.src_index = n_src++;
      }
      dec_.push_back(ly);
    }
  }
  // K / V projections of the memory for every cross-attention layer as ONE contraction
  RT_OK(make_linear_from(kvw, n_src * 2 * d_, d_, 1, &kvb, &dec_kv_));
  RT_OK(make_norm("decoder.after_norm.", &dec_after_));
  RT_OK(make_linear("decoder.output_layer.weight", "decoder.output_layer.bias", &dec_out_));
  vocab_ = dec_out_.n;
  host_.clear();  // the fp32 host copies are no longer needed
  // activation slab: sized for one batch of max_rows encoder rows (decoder rows <= encoder rows)
  const size_t r = (size_t)max_rows_ + 64;
  size_t cap = r * lfr_n_ * n_mels_ * 4 + r * din_ * 4  // fbank, feats
               + r * (size_t)np_ * (kp64(din_) + 7 * kp64(d_) + 2 * kp64(ff_) + kp64(3 * d_) + kp64(dec_kv_.n)) * 2  // splits
               + r * (size_t)d_ * 4 * 16 + r * (size_t)ff_ * 4 + r * (size_t)((vocab_ + 3) / 4 * 4) * 4 + (64u << 20);
  if (has_stamps_)  // upsampled rows: split of the conv output, xg [3r, 8H], the LSTM kernel's partial sums, alphas and fires
    cap += r * (size_t)np_ * 3 * 512 * 2 + 3 * r * (size_t)(8 * 512 + ts_parts_ + 4) * 4 + (8u << 20);
  RT_CUDA(cudaMalloc(&arena_.base, cap));
  arena_.cap = cap;
  RT_CUDA(cudaStreamSynchronize(stream_));
  return FB200_OK;
}

// ---------------------------------------------------------------- one packed batch --------
int Recognizer::reserve_wave(size_t samples) {
  if (samples <= wave_cap_) return FB200_OK;
  RT_CUDA(cudaStreamSynchronize(copy_stream_));
  RT_CUDA(cudaStreamSynchronize(stream_));
  for (auto& g : graphs_) cudaGraphExecDestroy(g.second);  // the captured launches read the old waveform buffers
  graphs_.clear();
  for (int i = 0; i < 2; ++i) {
    if (dwave_[i]) cudaFree(dwave_[i]);
    dwave_[i] = nullptr;
    RT_CUDA(cudaMalloc(&dwave_[i], samples * sizeof(float)));
  }
  wave_cap_ = samples;
  return FB200_OK;
}

// Waveforms of one batch: caller memory -> device slot, on the copy stream.
int Recognizer::stage(const std::vector<int>& batch, int slot, const float* const* waves, const int* lens) {
  size_t off = 0;
  for (int i : batch) {
    RT_CUDA(cudaMemcpyAsync(dwave_[slot] + off, waves[i], (size_t)lens[i] * sizeof(float), cudaMemcpyHostToDevice, copy_stream_));
    off += (size_t)lens[i];
  }
  RT_CUDA(cudaEventRecord(copied_[slot], copy_stream_));
  return FB200_OK;
}

static double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

template <typename F>
int Recognizer::run_graphed(const GraphKey& key, bool allowed, F&& issue) {
  if (!allowed || !use_graphs_) return issue();
  auto it = graphs_.find(key);
  if (it != graphs_.end()) {
    RT_CUDA(cudaGraphLaunch(it->second, stream_));
    return FB200_OK;
  }
  if (++graph_seen_[key] < 2 || graphs_.size() >= 64) return issue();
  if (cudaStreamBeginCapture(stream_, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
    cudaGetLastError();
    return issue();
  }
  const int rc = issue();
  cudaGraph_t graph = nullptr;
  const cudaError_t e = cudaStreamEndCapture(stream_, &graph);
  cudaGraphExec_t exec = nullptr;
  if (rc != FB200_OK || e != cudaSuccess || graph == nullptr || cudaGraphInstantiate(&exec, graph, 0) != cudaSuccess) {
    if (graph) cudaGraphDestroy(graph);
    cudaGetLastError();
    use_graphs_ = false;  // something on the stretch cannot be captured here: stay eager from now on
    return rc != FB200_OK ? rc : issue();
  }
  cudaGraphDestroy(graph);
  graphs_[key] = exec;
  RT_CUDA(cudaGraphLaunch(exec, stream_));
  return FB200_OK;
}

// The 50 SAN-M encoder layers + after_norm over packed rows (the per-op calls; inside fb200_stack_begin / _end they are
// recorded instead of launched).  b.encs != nullptr: after_norm also as a plain split (the decoder's memory operand).
int Recognizer::encoder_ops(const EncBufs& b, int R, int B, int max_t, void* st) {
  const int np = np_, d = d_;
  const int wqkv = np * kp64(3 * d);
  const float scale = 1.f / sqrtf((float)(d / heads_));
  const float* x = b.feats;
  int ldx = din_;
  for (const EncLayer& ly : enc_) {
    bf16* a = ly.din == d ? b.xs : b.xs0;
    RT_OK(fb200_layernorm(x, ldx, R, ly.din, ly.norm1.g, ly.norm1.b, eps_, nullptr, 0, a, np, 0, nullptr, 0, st));
    RT_OK(fb200_gemm(a, R, np, ly.qkv.w, np, R, 3 * d, ly.din, 1, ly.qkv.bias, 0, nullptr, 0, b.v, d, 2 * d, b.qkv, np, 0, st));
    RT_OK(fb200_fsmn(b.v, d, b.d_cur, B, R, d, ly.fsmn, ly.fsmn_k, ly.din == d ? x : nullptr, ldx, b.mem, d, st));
    RT_OK(fb200_attention(b.qkv, wqkv, 0, wqkv / np, b.qkv, wqkv, d, b.qkv, wqkv, 2 * d, wqkv / np, np, b.d_cur, b.d_cur, b.plan, B, R, R,
                          max_t, heads_, d / heads_, scale, nullptr, 0, b.xs, np, 0, st));
    RT_OK(fb200_gemm(b.xs, R, np, ly.out.w, np, R, d, d, 1, ly.out.bias, 0, b.mem, d, b.x1, d, 0, nullptr, np, 0, st));
    RT_OK(fb200_layernorm(b.x1, d, R, d, ly.norm2.g, ly.norm2.b, eps_, nullptr, 0, b.xs, np, 0, nullptr, 0, st));
    RT_OK(fb200_gemm(b.xs, R, np, ly.w1.w, np, R, ff_, d, 1, ly.w1.bias, 1, nullptr, 0, nullptr, 0, 0, b.h, np, 0, st));
    RT_OK(fb200_gemm(b.h, R, np, ly.w2.w, np, R, d, ff_, 1, ly.w2.bias, 0, b.x1, d, b.x0, d, 0, nullptr, np, 0, st));
    x = b.x0;
    ldx = d;
  }
  RT_OK(fb200_layernorm(x, d, R, d, enc_after_.g, enc_after_.b, eps_, b.enc, d, b.halo, np, 1, b.d_cur, B, st));
  if (b.encs != nullptr) RT_OK(fb200_layernorm(x, d, R, d, enc_after_.g, enc_after_.b, eps_, nullptr, 0, b.encs, np, 0, nullptr, 0, st));
  return FB200_OK;
}

// K / V projection of the memory, the 16 decoder layers, after_norm and the vocabulary projection over packed token rows.
int Recognizer::decoder_ops(const DecBufs& b, int R, int L, int B, int max_l, int ldv, void* st) {
  const int np = np_, d = d_;
  const int wd = np * kp64(d), wkv = np * kp64(dec_kv_.n);
  const float scale = 1.f / sqrtf((float)(d / heads_));
  RT_OK(fb200_gemm(b.encs, R, np, dec_kv_.w, np, R, dec_kv_.n, d, 1, dec_kv_.bias, 0, nullptr, 0, nullptr, 0, 0, b.kv, np, 0, st));
  const float* y = b.embeds;
  for (const DecLayer& ly : dec_) {
    RT_OK(fb200_layernorm(y, d, L, d, ly.norm1.g, ly.norm1.b, eps_, nullptr, 0, b.ys, np, 0, nullptr, 0, st));
    RT_OK(fb200_gemm(b.ys, L, np, ly.w1.w, np, L, ff_, d, 1, ly.w1.bias, 1, nullptr, 0, b.hf, ff_, 0, nullptr, np, 0, st));
    RT_OK(fb200_layernorm(b.hf, ff_, L, ff_, ly.ffn_norm.g, ly.ffn_norm.b, eps_, nullptr, 0, b.hs, np, 0, nullptr, 0, st));
    float* tt = (y == b.t) ? b.t2 : b.t;
    RT_OK(fb200_gemm(b.hs, L, np, ly.w2.w, np, L, d, ff_, 1, ly.w2.bias, 0, nullptr, 0, tt, d, 0, nullptr, np, 0, st));
    if (!ly.has_fsmn) {  // decoders3: FFN only, no residual (paraformer/decoder.py:102-108)
      y = tt;
      continue;
    }
    RT_OK(fb200_layernorm(tt, d, L, d, ly.norm2.g, ly.norm2.b, eps_, b.n2, d, nullptr, np, 0, nullptr, 0, st));
    float* yo = (y == b.b0) ? b.b1 : b.b0;
    RT_OK(fb200_fsmn(b.n2, d, b.d_cut, B, L, d, ly.fsmn, ly.fsmn_k, y, d, yo, d, st));  // x + conv(n2) + n2
    if (ly.has_src) {
      RT_OK(fb200_layernorm(yo, d, L, d, ly.norm3.g, ly.norm3.b, eps_, nullptr, 0, b.ys, np, 0, nullptr, 0, st));
      RT_OK(fb200_gemm(b.ys, L, np, ly.q.w, np, L, d, d, 1, ly.q.bias, 0, nullptr, 0, nullptr, 0, 0, b.qs, np, 0, st));
      const int j = ly.src_index;
      RT_OK(fb200_attention(b.qs, wd, 0, wd / np, b.kv, wkv, j * 2 * d, b.kv, wkv, j * 2 * d + d, wkv / np, np, b.d_cut, b.d_cur, b.dplan,
                            B, L, R, max_l, heads_, d / heads_, scale, nullptr, 0, b.ys, np, 0, st));
      float* xo = (yo == b.b0) ? b.b1 : b.b0;
      RT_OK(fb200_gemm(b.ys, L, np, ly.out.w, np, L, d, d, 1, ly.out.bias, 0, yo, d, xo, d, 0, nullptr, np, 0, st));
      y = xo;
    } else {
      y = yo;
    }
  }
  RT_OK(fb200_layernorm(y, d, L, d, dec_after_.g, dec_after_.b, eps_, nullptr, 0, b.ys, np, 0, nullptr, 0, st));
  if (b.logits != nullptr)  // (recorded stack programs: the logits are written, the pick is a launch of its own)
    RT_OK(fb200_gemm(b.ys, L, np, dec_out_.w, np, L, vocab_, d, 1, dec_out_.bias, 0, nullptr, 0, b.logits, ldv, 0, nullptr, np, 0, st));
  return FB200_OK;
}

// Buffers (capacity-sized, zeroed once) and the two programs of the small-request path.
int Recognizer::small_init() {
  const int np = np_, d = d_, RC = kStackRows, LC = kStackTokens, BC = kStackBatch;
  const int wd = np * kp64(d), wdin = np * kp64(din_), wff = np * kp64(ff_), wqkv = np * kp64(3 * d), wkv = np * kp64(dec_kv_.n);
  const int ldv = (vocab_ + 3) / 4 * 4;
  auto dev = [&](size_t bytes) -> void* {
    void* p = nullptr;
    if (cudaMalloc(&p, bytes) != cudaSuccess || cudaMemset(p, 0, bytes) != cudaSuccess) return nullptr;
    owned_.push_back(p);
    return p;
  };
  EncBufs& e = small_enc_;
  DecBufs& c = small_dec_;
  e.feats = (float*)dev((size_t)RC * din_ * 4);
  e.xs0 = (bf16*)dev((size_t)RC * wdin * 2);
  e.xs = (bf16*)dev((size_t)RC * wd * 2);
  e.h = (bf16*)dev((size_t)RC * wff * 2);
  e.qkv = (bf16*)dev((size_t)RC * wqkv * 2);
  e.v = (float*)dev((size_t)RC * d * 4);
  e.mem = (float*)dev((size_t)RC * d * 4);
  e.x0 = (float*)dev((size_t)RC * d * 4);
  e.x1 = (float*)dev((size_t)RC * d * 4);
  e.enc = (float*)dev((size_t)RC * d * 4);
  e.halo = (bf16*)dev((size_t)(RC + BC + 1) * wd * 2);
  e.encs = (bf16*)dev((size_t)RC * wd * 2);
  e.d_cur = (int32_t*)dev((BC + 1) * 4);
  e.plan = nullptr;
  c.encs = e.encs;
  c.kv = (bf16*)dev((size_t)RC * wkv * 2);
  c.qs = (bf16*)dev((size_t)LC * wd * 2);
  c.ys = (bf16*)dev((size_t)LC * wd * 2);
  c.hs = (bf16*)dev((size_t)LC * wff * 2);
  c.hf = (float*)dev((size_t)LC * ff_ * 4);
  c.t = (float*)dev((size_t)LC * d * 4);
  c.t2 = (float*)dev((size_t)LC * d * 4);
  c.n2 = (float*)dev((size_t)LC * d * 4);
  c.b0 = (float*)dev((size_t)LC * d * 4);
  c.b1 = (float*)dev((size_t)LC * d * 4);
  c.embeds = (float*)dev((size_t)LC * d * 4);
  c.logits = (float*)dev((size_t)LC * ldv * 4);
  c.d_cut = (int32_t*)dev((BC + 1) * 4);
  c.d_cur = e.d_cur;
  c.dplan = nullptr;
  if (!e.feats || !e.xs0 || !e.xs || !e.h || !e.qkv || !e.v || !e.mem || !e.x0 || !e.x1 || !e.enc || !e.halo || !e.encs || !e.d_cur ||
      !c.kv || !c.qs || !c.ys || !c.hs || !c.hf || !c.t || !c.t2 || !c.n2 || !c.b0 || !c.b1 || !c.embeds || !c.logits || !c.d_cut) {
    set_error("runtime: device memory for the small-request buffers");
    return FB200_ECUDA;
  }
  const int32_t edims[2] = {RC, BC};
  RT_OK(fb200_stack_begin(&enc_prog_, edims, 2));
  int rc = encoder_ops(e, RC, BC, RC, stream_);
  if (rc != FB200_OK) {
    fb200_stack_destroy(enc_prog_);
    enc_prog_ = nullptr;
    return rc;
  }
  RT_OK(fb200_stack_end(enc_prog_));
  const int32_t ddims[3] = {LC, RC, BC};
  RT_OK(fb200_stack_begin(&dec_prog_, ddims, 3));
  rc = decoder_ops(c, RC, LC, BC, LC, ldv, stream_);
  if (rc != FB200_OK) {
    fb200_stack_destroy(dec_prog_);
    dec_prog_ = nullptr;
    return rc;
  }
  RT_OK(fb200_stack_end(dec_prog_));
  small_ready_ = true;
  return FB200_OK;
}

int Recognizer::forward(const std::vector<int>& batch, int slot, const int* lens, Result* res, const std::vector<int>* next,
                        const float* const* waves) {
  static const bool trace = getenv("FB200_RT_TRACE") != nullptr;  // host-side phase times of every batch on stderr
  const double t_begin = now_ms();
  double t_launched = 0, t_staged = 0, t_tokens = 0, t_dec = 0;
  // The next batch's waveforms are uploaded by a helper thread while this one issues the kernels of the current
  // batch: the issuing thread runs only as far ahead of the GPU as the launch queue allows, so an upload done here
  // after the last launch would leave the GPU idle for its whole duration (measured: 6.5 ms per batch, 90 ms per step).
  int stage_rc = FB200_OK;
  std::thread uploader;
  if (next != nullptr) uploader = std::thread([&]() {
    cudaSetDevice(device_);
    stage_rc = stage(*next, slot ^ 1, waves, lens);
  });
  struct Joiner {
    std::thread& t;
    ~Joiner() { if (t.joinable()) t.join(); }
  } joiner{uploader};
  const int B = (int)batch.size(), np = np_, d = d_;
  void* st = stream_;
  arena_.off = 0;
  std::vector<int64_t> cu_samples(B + 1, 0);
  std::vector<int32_t> cu_frames(B + 1, 0), cu_rows(B + 1, 0);
  int max_t = 0;
  for (int i = 0; i < B; ++i) {
    const int n = lens[batch[i]];
    const int frames = n < frame_len_ ? 0 : 1 + (n - frame_len_) / frame_shift_;
    const int rows = (frames + lfr_n_ - 1) / lfr_n_;
    cu_samples[i + 1] = cu_samples[i] + n;
    cu_frames[i + 1] = cu_frames[i] + frames;
    cu_rows[i + 1] = cu_rows[i] + rows;
    max_t = std::max(max_t, rows);
  }
  const int F = cu_frames[B], R = cu_rows[B];
  // small request: the encoder / decoder stretches run from the recorded programs over their own fixed buffers
  bool small = use_stack_ && np_ <= 2 && B <= kStackBatch && R > 0 && R <= kStackRows;
  if (small && !small_ready_) {
    static const bool off = getenv("FB200_STACK") != nullptr && atoi(getenv("FB200_STACK")) == 0;
    if (off || small_init() != FB200_OK) use_stack_ = small = false;
  }
  const float* wave = dwave_[slot];
  int64_t* d_cus = arena_.get<int64_t>(B + 1);
  int32_t *d_cuf = arena_.get<int32_t>(B + 1), *d_cur = arena_.get<int32_t>(B + 1), *d_cut = arena_.get<int32_t>(B + 1);
  float* fb = arena_.get<float>((size_t)F * n_mels_);
  float* feats = arena_.get<float>((size_t)R * din_);
  const int wd = np * kp64(d), wdin = np * kp64(din_), wff = np * kp64(ff_), wqkv = np * kp64(3 * d);
  bf16 *xs0 = arena_.get<bf16>((size_t)R * wdin), *xs = arena_.get<bf16>((size_t)R * wd), *h = arena_.get<bf16>((size_t)R * wff);
  bf16* qkv = arena_.get<bf16>((size_t)R * wqkv);
  float *v = arena_.get<float>((size_t)R * d), *mem = arena_.get<float>((size_t)R * d), *x0 = arena_.get<float>((size_t)R * d),
        *x1 = arena_.get<float>((size_t)R * d), *enc = arena_.get<float>((size_t)R * d);
  bf16* halo = arena_.get<bf16>((size_t)(R + B + 1) * wd);
  const int plan_cap = R / 128 + B;
  int32_t* plan = arena_.get<int32_t>(4 + 4 * (size_t)plan_cap);
  float *conv = arena_.get<float>((size_t)(R + B) * d), *alphas = arena_.get<float>(R), *fire_rem = arena_.get<float>(R + B);
  int32_t *token_num = arena_.get<int32_t>(B), *fire_pos = arena_.get<int32_t>(R + B);
  if (!arena_.ok()) {
    set_error("runtime: activation slab too small for a batch of %d rows", R);
    return FB200_EINVAL;
  }
  EncBufs eb = {feats, xs0, xs, h, qkv, v, mem, x0, x1, enc, halo, nullptr, d_cur, plan};
  if (small) {
    eb = small_enc_;
    feats = eb.feats;
    enc = eb.enc;
    halo = eb.halo;
    d_cur = eb.d_cur;
  }
  RT_CUDA(cudaStreamWaitEvent(stream_, copied_[slot], 0));  // this batch's waveforms are on the device
  RT_CUDA(cudaMemcpyAsync(d_cus, cu_samples.data(), (B + 1) * 8, cudaMemcpyHostToDevice, stream_));
  RT_CUDA(cudaMemcpyAsync(d_cuf, cu_frames.data(), (B + 1) * 4, cudaMemcpyHostToDevice, stream_));
  RT_CUDA(cudaMemcpyAsync(d_cur, cu_rows.data(), (B + 1) * 4, cudaMemcpyHostToDevice, stream_));
  const float scale = 1.f / sqrtf((float)(d / heads_));
  auto issue_encoder = [&]() -> int {
  // frontend + encoder prologue (x * sqrt(d) + PE)
  RT_OK(fb200_fbank(wave, d_cus, d_cuf, B, F, window_, frame_len_, frame_shift_, nfft_, mel_, n_mels_, 32768.f, 0.97f, fb, st));
  RT_OK(fb200_lfr_cmvn(fb, d_cuf, d_cur, B, R, n_mels_, lfr_m_, lfr_n_, shift_, scale_, pe_inv_, sqrtf((float)d), 0, feats, din_, st));
  // encoder
  if (small) {
    const int32_t dims[2] = {R, B};
    RT_OK(fb200_stack_run(enc_prog_, dims, 2, st));
  } else {
    RT_OK(fb200_attention_plan(d_cur, d_cur, B, plan, plan_cap, st));
    RT_OK(encoder_ops(eb, R, B, max_t, st));
  }
  // predictor: k = 3 conv over the halo layout, alpha head, integrate-and-fire scan
  RT_OK(fb200_gemm(halo, R + B + 1, np, cif_conv_.w, np, R + B - 1, d, d, 3, cif_conv_.bias, 1, nullptr, 0, conv, d, 0, nullptr, np, 0, st));
  RT_OK(fb200_cif_alpha(conv, d, d_cur, B, R, d, cif_w_, cif_b_, 1.f, 0.f, alphas, st));
  RT_OK(fb200_cif_scan(alphas, d_cur, B, tail_, token_num, fire_pos, fire_rem, nullptr, st));
  return FB200_OK;
  };
  RT_OK(run_graphed(GraphKey(slot, B, F, R, max_t, 0, 0), next == nullptr && !small, issue_encoder));
  std::vector<int32_t> tok(B), cu_tok(B + 1, 0);
  RT_CUDA(cudaMemcpyAsync(tok.data(), token_num, B * 4, cudaMemcpyDeviceToHost, stream_));
  t_launched = now_ms();
  if (uploader.joinable()) uploader.join();
  if (stage_rc != FB200_OK) {
    set_error("runtime: upload of the next batch failed (%d)", stage_rc);
    return stage_rc;
  }
  t_staged = now_ms();
  RT_CUDA(cudaStreamSynchronize(stream_));  // the one device -> host read before the decoder: [B] token counts
  t_tokens = now_ms();
  int max_l = 0;
  for (int i = 0; i < B; ++i) cu_tok[i + 1] = cu_tok[i] + tok[i], max_l = std::max(max_l, tok[i]);
  const int L = cu_tok[B];
  if (L == 0 || max_l < 1) return FB200_OK;  // no utterance has a token (paraformer/model.py:514-515)
  const bool small_dec = small && L <= kStackTokens;
  if (small_dec) d_cut = small_dec_.d_cut;
  RT_CUDA(cudaMemcpyAsync(d_cut, cu_tok.data(), (B + 1) * 4, cudaMemcpyHostToDevice, stream_));
  float* embeds = arena_.get<float>((size_t)L * d);
  // decoder
  const int wkv = np * kp64(dec_kv_.n);
  bf16 *encs = arena_.get<bf16>((size_t)R * wd), *kv = arena_.get<bf16>((size_t)R * wkv), *qs = arena_.get<bf16>((size_t)L * wd),
       *ys = arena_.get<bf16>((size_t)L * wd), *hs = arena_.get<bf16>((size_t)L * wff);
  float *hf = arena_.get<float>((size_t)L * ff_), *t = arena_.get<float>((size_t)L * d), *t2 = arena_.get<float>((size_t)L * d),
        *n2 = arena_.get<float>((size_t)L * d), *b0 = arena_.get<float>((size_t)L * d), *b1 = arena_.get<float>((size_t)L * d);
  const int dplan_cap = L / 128 + B;
  int32_t* dplan = arena_.get<int32_t>(4 + 4 * (size_t)dplan_cap);
  const int ldv = (vocab_ + 3) / 4 * 4;
  // per-op path: vocabulary projection fused with the greedy pick (fb200_gemm_argmax: the [L, V] logits are never written)
  long long pick_bytes = 0;
  fb200_gemm_argmax_workspace(L, vocab_, &pick_bytes);
  float* logits = small_dec ? nullptr : arena_.get<float>((size_t)(pick_bytes + 3) / 4);
  int32_t* ids = arena_.get<int32_t>(L);
  float* logp = arena_.get<float>(L);
  if (!arena_.ok()) {
    set_error("runtime: activation slab too small for %d tokens", L);
    return FB200_EINVAL;
  }
  DecBufs db = {encs, kv, qs, ys, hs, hf, t, t2, n2, b0, b1, embeds, logits, d_cut, d_cur, dplan};
  if (small_dec) {
    db = small_dec_;
    embeds = db.embeds;
    logits = db.logits;
    encs = db.encs;  // (the time-stamp head reads the split of the encoder output too)
  }
  auto issue_decoder = [&]() -> int {  // CIF integration + decoder + greedy pick: ~110 launches, graph-replayed like the encoder
  RT_OK(fb200_cif_integrate(enc, d, alphas, d_cur, d_cut, B, L, d, tail_, fire_pos, fire_rem, embeds, d, st));
  if (small_dec) {
    const int32_t dims[3] = {L, R, B};
    RT_OK(fb200_stack_run(dec_prog_, dims, 3, st));
  } else {
    RT_OK(fb200_split_rows(enc, d, R, d, encs, np, st));
    RT_OK(fb200_attention_plan(d_cut, d_cur, B, dplan, dplan_cap, st));
    DecBufs nb = db;
    nb.logits = nullptr;
    RT_OK(decoder_ops(nb, R, L, B, max_l, ldv, st));
    RT_OK(fb200_gemm_argmax(db.ys, L, np, dec_out_.w, np, L, vocab_, d, dec_out_.bias, logits, pick_bytes, ids, logp, st));
    return FB200_OK;
  }
  RT_OK(fb200_argmax_logprob(logits, ldv, L, vocab_, ids, logp, st));
  return FB200_OK;
  };
  RT_OK(run_graphed(GraphKey(slot | 0x100, B, F, R, max_t, L, max_l), next == nullptr && !small_dec, issue_decoder));
  std::vector<int32_t> h_ids(L);
  RT_CUDA(cudaMemcpyAsync(h_ids.data(), ids, (size_t)L * 4, cudaMemcpyDeviceToHost, stream_));
  std::vector<int32_t> h_fire, h_fire_count, h_chars;
  if (has_stamps_) {
    // CifPredictorV3.get_upsample_timestamp + the fire search of ts_prediction_lfr6_standard (see funasr_b200/bicif.py)
    const int H = 512, up = 3, R3 = up * R;
    bf16* ups = arena_.get<bf16>((size_t)R * np * up * H);
    float* xg = arena_.get<float>((size_t)R3 * 8 * H);
    float* part = arena_.get<float>((size_t)ts_parts_ * R3);
    long long ws_bytes = 0;
    RT_OK(fb200_lstm_workspace_bytes(B, H, &ws_bytes));
    float* ws = arena_.get<float>((size_t)(ws_bytes + 3) / 4);
    float* us_alphas = arena_.get<float>(R3);
    int32_t *fire_idx = arena_.get<int32_t>(R3), *fire_count = arena_.get<int32_t>(B), *num_chars = arena_.get<int32_t>(B);
    if (!arena_.ok()) {
      set_error("runtime: activation slab too small for the time-stamp head of %d rows", R);
      return FB200_EINVAL;
    }
    RT_OK(fb200_gemm(encs, R, np, ts_up_.w, np, R, up * H, d, 1, ts_up_.bias, 0, nullptr, 0, nullptr, 0, 0, ups, np, 0, st));
    const int zero_tap = 0;
    for (int j = 0; j < up; ++j)  // sub-frame j of every encoder frame -> rows 3t + j of xg
      RT_OK(fb200_gemm_ex(ups, R, np * up * H, up * H, j * H, np, ts_proj_.w, np, R, 8 * H, H, 1, &zero_tap, &zero_tap, ts_proj_.bias, 0,
                          nullptr, 0, xg + (size_t)j * 8 * H, up * 8 * H, 0, nullptr, 1, 0, st));
    RT_OK(fb200_lstm_bidir(xg, 8 * H, d_cur, up, B, R3, up * max_t, H, ts_whh_, ts_wout_, ws, ws_bytes, part, nullptr, 0, st));
    const int32_t drop[3] = {0, 1, 2};
    RT_OK(fb200_us_fires(part, ts_parts_, R3, ts_bout_, 0.25f, 0.01f, (float)(1.0 - 1e-4), d_cur, up, B, token_num, ids, d_cut, drop,
                         us_alphas, nullptr, fire_idx, fire_count, num_chars, st));
    h_fire.resize(R3);
    h_fire_count.resize(B);
    h_chars.resize(B);
    RT_CUDA(cudaMemcpyAsync(h_fire.data(), fire_idx, (size_t)R3 * 4, cudaMemcpyDeviceToHost, stream_));
    RT_CUDA(cudaMemcpyAsync(h_fire_count.data(), fire_count, (size_t)B * 4, cudaMemcpyDeviceToHost, stream_));
    RT_CUDA(cudaMemcpyAsync(h_chars.data(), num_chars, (size_t)B * 4, cudaMemcpyDeviceToHost, stream_));
  }
  t_dec = now_ms();
  RT_CUDA(cudaStreamSynchronize(stream_));
  if (has_stamps_)
    for (int i = 0; i < B; ++i)
      if (tok[i] > 0)
        stamps_from_fires(h_fire.data() + 3 * (size_t)cu_rows[i], h_fire_count[i], 3 * (cu_rows[i + 1] - cu_rows[i]), h_chars[i],
                          &res->stamps[batch[i]]);
  if (trace)
    fprintf(stderr, "fb200 rt: batch %3d utts %6d rows %6d tok | launch enc %.2f  join upload %.2f  wait tokens %.2f  launch dec %.2f  wait ids %.2f ms\n",
            B, R, L, t_launched - t_begin, t_staged - t_launched, t_tokens - t_staged, t_dec - t_tokens, now_ms() - t_dec);
  for (int i = 0; i < B; ++i) {
    std::vector<int32_t>& out = res->ids[batch[i]];
    for (int k = cu_tok[i]; k < cu_tok[i + 1]; ++k)
      if (h_ids[k] > 2) out.push_back(h_ids[k]);  // blank, sos, eos dropped (paraformer/model.py:561-565)
  }
  return FB200_OK;
}

int Recognizer::recognize(const float* const* waves, const int* lens, int n, Result* res) {
  RT_CUDA(cudaSetDevice(device_));
  res->ids.assign(n, std::vector<int32_t>());
  res->text.assign(n, std::string());
  res->stamps.assign(n, std::vector<int32_t>());
  // length-sorted batches under the packed-row budget (ref. analogue: funasr/auto/auto_model.py:423-459)
  std::vector<int> rows(n), order;
  for (int i = 0; i < n; ++i) {
    const int frames = lens[i] < frame_len_ ? 0 : 1 + (lens[i] - frame_len_) / frame_shift_;
    rows[i] = (frames + lfr_n_ - 1) / lfr_n_;
    if (rows[i] >= 1) order.push_back(i);
  }
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return rows[a] > rows[b]; });
  std::vector<std::vector<int>> batches;
  std::vector<int> cur;
  int cur_rows = 0;
  for (int i : order) {
    if (rows[i] > max_rows_) {
      set_error("runtime: an utterance of %d encoder rows exceeds max_rows %d", rows[i], max_rows_);
      return FB200_EINVAL;
    }
    // the time-stamp head's LSTM launch takes at most kMaxStampBatch utterances (fb200_lstm_bidir: batch <= 1024)
    if (!cur.empty() && (cur_rows + rows[i] > max_rows_ || (has_stamps_ && (int)cur.size() >= kMaxStampBatch))) {
      batches.push_back(cur);
      cur.clear();
      cur_rows = 0;
    }
    cur.push_back(i);
    cur_rows += rows[i];
  }
  if (!cur.empty()) batches.push_back(cur);
  if (batches.size() >= 2) {
    // nothing overlaps the upload of the first batch: cut it at a third of the row budget, the rest arrives under its kernels
    size_t k = 0;
    int acc = 0;
    while (k < batches[0].size() && (k == 0 || acc + rows[batches[0][k]] <= max_rows_ / 3)) acc += rows[batches[0][k++]];
    if (k < batches[0].size()) {
      std::vector<int> rest(batches[0].begin() + (long)k, batches[0].end());
      batches[0].resize(k);
      batches.insert(batches.begin() + 1, rest);
    }
  }
  size_t need = 0;
  for (const auto& b : batches) {
    size_t smp = 0;
    for (int i : b) smp += (size_t)lens[i];
    need = std::max(need, smp);
  }
  if (batches.empty()) return FB200_OK;
  RT_OK(reserve_wave(need));
  RT_OK(stage(batches[0], 0, waves, lens));
  for (size_t k = 0; k < batches.size(); ++k)
    RT_OK(forward(batches[k], (int)(k & 1), lens, res, k + 1 < batches.size() ? &batches[k + 1] : nullptr, waves));
  for (int i = 0; i < n; ++i) {
    std::string& s = res->text[i];
    for (int32_t id : res->ids[i]) {
      if (!s.empty() && tokens_.empty()) s.push_back(' ');
      if (tokens_.empty()) s += std::to_string(id);
      else if ((size_t)id < tokens_.size()) s += tokens_[id];
    }
  }
  return FB200_OK;
}

}  // namespace rt
}  // namespace fb200

using fb200::rt::Recognizer;
using fb200::rt::Result;
#define RT_API extern "C" __attribute__((visibility("default")))

RT_API void* fb200_offline_init(const char* model_dir, int device, int precision, int max_rows) {
  if (model_dir == nullptr || (precision != 0 && precision != 1)) {
    fb200::set_error("offline_init: model_dir / precision (0 = fp32 parity, 1 = bf16)");
    return nullptr;
  }
  Recognizer* r = nullptr;
  try {
    r = new Recognizer();
    if (r->init(model_dir, device, precision == 0 ? 2 : 1, max_rows > 0 ? max_rows : 56832) != FB200_OK) {
      delete r;
      return nullptr;
    }
  } catch (const std::exception& ex) {  // nothing may unwind through the C boundary
    fb200::set_error("offline_init: %s", ex.what());
    delete r;
    return nullptr;
  }
  return r;
}

RT_API void* fb200_offline_infer(void* handle, const float* const* waves, const int32_t* lens, int count) {
  if (handle == nullptr || count < 0 || (count > 0 && (waves == nullptr || lens == nullptr))) {
    fb200::set_error("offline_infer: bad arguments");
    return nullptr;
  }
  for (int i = 0; i < count; ++i)
    if (lens[i] < 0 || (lens[i] > 0 && waves[i] == nullptr)) {
      fb200::set_error("offline_infer: utterance %d has length %d / a null waveform", i, lens[i]);
      return nullptr;
    }
  Result* res = nullptr;
  try {
    res = new Result();
    if (static_cast<Recognizer*>(handle)->recognize(waves, lens, count, res) != FB200_OK) {
      delete res;
      return nullptr;
    }
  } catch (const std::exception& ex) {
    fb200::set_error("offline_infer: %s", ex.what());
    delete res;
    return nullptr;
  }
  return res;
}

RT_API void* fb200_offline_infer_buffer(void* handle, const char* pcm16, int n_bytes, int sampling_rate) {
  if (sampling_rate != 16000 || pcm16 == nullptr || n_bytes < 0) {
    fb200::set_error("offline_infer_buffer: 16 kHz mono PCM16 only (got %d Hz)", sampling_rate);
    return nullptr;
  }
  const int n = n_bytes / 2;
  std::vector<float> wave;
  try {
    wave.resize(n);
  } catch (const std::exception& ex) {
    fb200::set_error("offline_infer_buffer: %s", ex.what());
    return nullptr;
  }
  const int16_t* s = reinterpret_cast<const int16_t*>(pcm16);
  for (int i = 0; i < n; ++i) wave[i] = (float)s[i] / 32768.f;  // the frontend scales by 1 << 15 again (wav_frontend.py:130)
  const float* w = wave.data();
  const int32_t len = n;
  return fb200_offline_infer(handle, &w, &len, 1);
}

RT_API int fb200_result_count(void* result) { return result ? (int)static_cast<Result*>(result)->ids.size() : 0; }
RT_API const char* fb200_result_text(void* result, int index) {
  Result* r = static_cast<Result*>(result);
  return (r && index >= 0 && index < (int)r->text.size()) ? r->text[index].c_str() : nullptr;
}
RT_API const int32_t* fb200_result_ids(void* result, int index, int* count) {
  Result* r = static_cast<Result*>(result);
  if (!r || index < 0 || index >= (int)r->ids.size()) {
    if (count) *count = 0;
    return nullptr;
  }
  if (count) *count = (int)r->ids[index].size();
  return r->ids[index].data();
}
RT_API const int32_t* fb200_result_stamps(void* result, int index, int* count) {
  Result* r = static_cast<Result*>(result);
  if (!r || index < 0 || index >= (int)r->stamps.size()) {
    if (count) *count = 0;
    return nullptr;
  }
  if (count) *count = (int)r->stamps[index].size() / 2;
  return r->stamps[index].data();
}
RT_API void fb200_free_result(void* result) { delete static_cast<Result*>(result); }
RT_API void fb200_offline_uninit(void* handle) { delete static_cast<Recognizer*>(handle); }
