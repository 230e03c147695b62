// Stand-alone caller of the native offline runtime (no Python, no torch): the shape of the reference's
// runtime/onnxruntime/bin/funasr-onnx-offline.cpp over funasrruntime.h, here over include/funasr_b200.h.
//
//   g++ -O2 -std=c++17 -Iinclude examples/offline_recognize.cpp -Lfunasr_b200 -lfunasr_b200 -Wl,-rpath,$PWD/funasr_b200 -o examples/offline_recognize
//   examples/offline_recognize <model_dir> <a.pcm|a.wav> [b.pcm ...]      (16 kHz mono PCM16; a 44-byte RIFF header is skipped)
//
// All files are decoded as ONE packed batch; prints "<file>\t<token ids>\t<text>" per input.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <fstream>
#include <string>
#include <vector>

#include "funasr_b200.h"

static bool read_pcm16(const std::string& path, std::vector<float>* out) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  std::vector<char> raw((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
  size_t off = (raw.size() >= 44 && memcmp(raw.data(), "RIFF", 4) == 0) ? 44 : 0;
  const size_t n = (raw.size() - off) / 2;
  out->resize(n);
  const int16_t* s = reinterpret_cast<const int16_t*>(raw.data() + off);
  for (size_t i = 0; i < n; ++i) (*out)[i] = (float)s[i] / 32768.f;
  return true;
}

int main(int argc, char** argv) {
  if (argc < 3) {
    fprintf(stderr, "usage: %s <model_dir> <audio.pcm|audio.wav> [...]\n", argv[0]);
    return 2;
  }
  void* h = fb200_offline_init(argv[1], 0, 0, 0);
  if (h == nullptr) {
    fprintf(stderr, "fb200_offline_init: %s\n", fb200_last_error());
    return 1;
  }
  std::vector<std::vector<float>> waves(argc - 2);
  std::vector<const float*> ptrs;
  std::vector<int32_t> lens;
  for (int i = 2; i < argc; ++i) {
    if (!read_pcm16(argv[i], &waves[i - 2])) {
      fprintf(stderr, "cannot read %s\n", argv[i]);
      return 1;
    }
    ptrs.push_back(waves[i - 2].data());
    lens.push_back((int32_t)waves[i - 2].size());
  }
  void* res = fb200_offline_infer(h, ptrs.data(), lens.data(), (int)ptrs.size());
  if (res == nullptr) {
    fprintf(stderr, "fb200_offline_infer: %s\n", fb200_last_error());
    return 1;
  }
  for (int i = 0; i < fb200_result_count(res); ++i) {
    int n = 0;
    const int32_t* ids = fb200_result_ids(res, i, &n);
    printf("%s\t", argv[i + 2]);
    for (int k = 0; k < n; ++k) printf(k ? " %d" : "%d", ids[k]);
    printf("\t%s", fb200_result_text(res, i));
    int ns = 0;
    const int32_t* st = fb200_result_stamps(res, i, &ns);  // [begin ms, end ms] per token (BiCifParaformer models)
    for (int k = 0; k < ns; ++k) printf(k ? " %d-%d" : "\t%d-%d", st[2 * k], st[2 * k + 1]);
    printf("\n");
  }
  fb200_free_result(res);
  fb200_offline_uninit(h);
  return 0;
}
