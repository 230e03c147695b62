// libfunasr_b200_compat.so -- the reference's C++ offline-stream API (runtime/onnxruntime/include/funasrruntime.h:69-74,
// 99-114) over this library's native runtime (fb200_offline_*, csrc/runtime.cu), with the reference's names AND C++
// signatures (std::map / std::string / std::vector arguments: the symbols are C++-mangled, as the reference's are), so a
// server written against funasrruntime.h -- runtime/websocket/bin/websocket-server.cpp:81-88 calls FunOfflineInferBuffer,
// FunASRGetResult, FunASRGetStamp, FunASRGetStampSents, FunASRFreeResult -- links against this file's library unchanged.
// tests/test_compat.py compiles a caller against the reference's own header and links it here.
//
//   reference (src/funasrruntime.cpp)                       here
//   FunOfflineInit(model_path, thread_num, use_gpu, batch)  model_path["model-dir"] holds model.fb200 (+ am.mvn, tokens.json);
//                                                           "quantize" == "true" selects the bf16 throughput mode; GPU only
//   FunOfflineInferBuffer(h, buf, n, mode, cb, hw_emb, fs,  16 kHz mono PCM16 ("pcm") -> one result; VAD / punctuation / ITN /
//       wav_format, itn, dec_handle, svs_lang, svs_itn)     hot-word embeddings / WFST decoding are other models: not on this path
//   FunOfflineInfer(h, file, ...)                           .pcm and PCM16 .wav files
//   FunASRGetResult / GetStamp / GetStampSents /            text; "[[begin ms,end ms],...]" (:296-308) when the model directory
//   GetRetNumber / GetRetSnippetTime / FreeResult           has the CifPredictorV3 head, else ""; ""; 1; seconds of audio
//   FunOfflineReset / FunOfflineUninit
#include <stdint.h>
#include <string.h>

#include <fstream>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/funasr_b200.h"

#define FB_COMPAT_API __attribute__((visibility("default")))

typedef void* FUNASR_HANDLE;
typedef void* FUNASR_RESULT;
typedef void* FUNASR_DEC_HANDLE;
typedef enum { RASR_NONE = -1, RASRM_CTC_GREEDY_SEARCH = 0, RASRM_CTC_RPEFIX_BEAM_SEARCH = 1, RASRM_ATTENSION_RESCORING = 2 } FUNASR_MODE;
typedef enum { ASR_OFFLINE = 0, ASR_ONLINE = 1, ASR_TWO_PASS = 2 } ASR_TYPE;
typedef void (*QM_CALLBACK)(int cur_step, int n_total);

namespace {

struct CompatResult {  // FUNASR_RECOG_RESULT's fields that the getters expose
  std::string msg, stamp, stamp_sents;
  float snippet_time = 0.f;
};

struct CompatHandle {
  void* rt = nullptr;
};

CompatResult* run(CompatHandle* h, const char* pcm16, int n_bytes, int sampling_rate) {
  CompatResult* res = new CompatResult();
  res->snippet_time = (float)(n_bytes / 2) / (float)(sampling_rate > 0 ? sampling_rate : 16000);
  if (n_bytes < 2) return res;  // the reference returns an empty result for empty audio (:233-236)
  void* r = fb200_offline_infer_buffer(h->rt, pcm16, n_bytes, sampling_rate);
  if (r == nullptr) {
    delete res;
    return nullptr;
  }
  if (fb200_result_count(r) > 0) {
    const char* text = fb200_result_text(r, 0);
    if (text) res->msg = text;
    int n = 0;
    const int32_t* st = fb200_result_stamps(r, 0, &n);
    if (st != nullptr && n > 0) {
      res->stamp = "[";
      for (int i = 0; i < n; ++i) {
        if (i) res->stamp += ",";
        res->stamp += "[" + std::to_string(st[2 * i]) + "," + std::to_string(st[2 * i + 1]) + "]";
      }
      res->stamp += "]";
    }
  }
  fb200_free_result(r);
  return res;
}

// PCM16 samples of a RIFF/WAVE file (first "data" chunk; format 1, mono) or of a raw .pcm file
bool load_pcm16(const char* path, std::vector<char>* out, int* sampling_rate) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  std::vector<char> raw((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
  if (raw.size() >= 12 && memcmp(raw.data(), "RIFF", 4) == 0 && memcmp(raw.data() + 8, "WAVE", 4) == 0) {
    size_t o = 12;
    bool fmt_ok = false;
    while (o + 8 <= raw.size()) {
      uint32_t len = 0;
      memcpy(&len, raw.data() + o + 4, 4);
      if (memcmp(raw.data() + o, "fmt ", 4) == 0 && len >= 16 && o + 8 + 16 <= raw.size()) {
        uint16_t format = 0, channels = 0, bits = 0;
        uint32_t rate = 0;
        memcpy(&format, raw.data() + o + 8, 2);
        memcpy(&channels, raw.data() + o + 10, 2);
        memcpy(&rate, raw.data() + o + 12, 4);
        memcpy(&bits, raw.data() + o + 22, 2);
        fmt_ok = format == 1 && channels == 1 && bits == 16;
        *sampling_rate = (int)rate;
      } else if (memcmp(raw.data() + o, "data", 4) == 0) {
        if (!fmt_ok) return false;
        const size_t n = std::min<size_t>(len, raw.size() - o - 8);
        out->assign(raw.begin() + o + 8, raw.begin() + o + 8 + n);
        return true;
      }
      o += 8 + (size_t)len + (len & 1u);
    }
    return false;
  }
  out->swap(raw);
  return true;
}

}  // namespace

FB_COMPAT_API FUNASR_HANDLE FunOfflineInit(std::map<std::string, std::string>& model_path, int thread_num, bool use_gpu, int batch_size) {
  (void)thread_num;
  (void)use_gpu;     // this back-end runs on the GPU only (no CPU fallback)
  (void)batch_size;  // requests are packed by the runtime itself
  auto it = model_path.find("model-dir");
  if (it == model_path.end()) return nullptr;
  auto q = model_path.find("quantize");
  const int precision = (q != model_path.end() && (q->second == "true" || q->second == "True" || q->second == "1")) ? 1 : 0;
  int device = 0;
  auto d = model_path.find("gpu-id");
  if (d != model_path.end()) device = atoi(d->second.c_str());
  void* rt = fb200_offline_init(it->second.c_str(), device, precision, 0);
  if (rt == nullptr) return nullptr;
  CompatHandle* h = new CompatHandle();
  h->rt = rt;
  return h;
}

FB_COMPAT_API void FunOfflineReset(FUNASR_HANDLE handle, FUNASR_DEC_HANDLE dec_handle) {
  (void)handle;
  (void)dec_handle;
}

FB_COMPAT_API FUNASR_RESULT FunOfflineInferBuffer(FUNASR_HANDLE handle, const char* sz_buf, int n_len, FUNASR_MODE mode, QM_CALLBACK fn_callback,
                                                  const std::vector<std::vector<float>>& hw_emb, int sampling_rate, std::string wav_format,
                                                  bool itn, FUNASR_DEC_HANDLE dec_handle, std::string svs_lang, bool svs_itn) {
  (void)mode; (void)fn_callback; (void)hw_emb; (void)itn; (void)dec_handle; (void)svs_lang; (void)svs_itn;
  CompatHandle* h = static_cast<CompatHandle*>(handle);
  if (h == nullptr || sz_buf == nullptr || n_len < 0) return nullptr;
  if (wav_format != "pcm" && wav_format != "PCM") return nullptr;  // other containers go through ffmpeg in the reference (:223)
  return run(h, sz_buf, n_len, sampling_rate);
}

FB_COMPAT_API FUNASR_RESULT FunOfflineInfer(FUNASR_HANDLE handle, const char* sz_filename, FUNASR_MODE mode, QM_CALLBACK fn_callback,
                                            const std::vector<std::vector<float>>& hw_emb, int sampling_rate, bool itn,
                                            FUNASR_DEC_HANDLE dec_handle) {
  (void)mode; (void)fn_callback; (void)hw_emb; (void)itn; (void)dec_handle;
  CompatHandle* h = static_cast<CompatHandle*>(handle);
  if (h == nullptr || sz_filename == nullptr) return nullptr;
  std::vector<char> pcm;
  int rate = sampling_rate;
  if (!load_pcm16(sz_filename, &pcm, &rate)) return nullptr;
  return run(h, pcm.data(), (int)pcm.size(), rate);
}

FB_COMPAT_API const std::vector<std::vector<float>> CompileHotwordEmbedding(FUNASR_HANDLE handle, std::string& hotwords, ASR_TYPE mode) {
  (void)handle; (void)hotwords; (void)mode;
  return {{0.0f}};  // the reference's "no hot words" value (funasrruntime.h: hw_emb={{0.0}}); hot words are the SeACo engine's job
}

FB_COMPAT_API void FunOfflineUninit(FUNASR_HANDLE handle) {
  CompatHandle* h = static_cast<CompatHandle*>(handle);
  if (h == nullptr) return;
  fb200_offline_uninit(h->rt);
  delete h;
}

FB_COMPAT_API const char* FunASRGetResult(FUNASR_RESULT result, int n_index) {
  (void)n_index;
  return result ? static_cast<CompatResult*>(result)->msg.c_str() : nullptr;
}
FB_COMPAT_API const char* FunASRGetStamp(FUNASR_RESULT result) { return result ? static_cast<CompatResult*>(result)->stamp.c_str() : nullptr; }
FB_COMPAT_API const char* FunASRGetStampSents(FUNASR_RESULT result) {
  return result ? static_cast<CompatResult*>(result)->stamp_sents.c_str() : nullptr;
}
FB_COMPAT_API const int FunASRGetRetNumber(FUNASR_RESULT result) { return result ? 1 : 0; }
FB_COMPAT_API const float FunASRGetRetSnippetTime(FUNASR_RESULT result) { return result ? static_cast<CompatResult*>(result)->snippet_time : 0.f; }
FB_COMPAT_API void FunASRFreeResult(FUNASR_RESULT result) { delete static_cast<CompatResult*>(result); }
