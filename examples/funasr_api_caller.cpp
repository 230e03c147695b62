// A caller written against the REFERENCE's header (runtime/onnxruntime/include/funasrruntime.h, unmodified -- passed with
// -I) that links libfunasr_b200_compat.so: the call sequence of runtime/websocket/bin/websocket-server.cpp:81-95
// (FunOfflineInferBuffer -> FunASRGetResult / FunASRGetStamp / FunASRGetStampSents -> FunASRFreeResult) around
// FunOfflineInit / FunOfflineUninit.
//
//   funasr_api_caller <model-dir> <file.pcm | file.wav> [quantize]
// prints "text\t<result>", "stamp\t<[[b,e],...]>", "seconds\t<audio length>".
#include <stdio.h>

#include <fstream>
#include <iterator>
#include <string>

#include "funasrruntime.h"

int main(int argc, char** argv) {
  if (argc < 3) {
    fprintf(stderr, "usage: %s <model-dir> <file.pcm|file.wav> [quantize]\n", argv[0]);
    return 2;
  }
  std::map<std::string, std::string> model_path;
  model_path["model-dir"] = argv[1];
  model_path["quantize"] = argc > 3 ? "true" : "false";
  FUNASR_HANDLE asr = FunOfflineInit(model_path, 1, true, 1);
  if (!asr) {
    fprintf(stderr, "FunOfflineInit failed\n");
    return 1;
  }
  std::vector<std::vector<float>> hw_emb = {{0.0f}};
  const std::string path = argv[2];
  FUNASR_RESULT result = nullptr;
  if (path.size() > 4 && path.substr(path.size() - 4) == ".wav") {
    result = FunOfflineInfer(asr, path.c_str(), RASR_NONE, nullptr, hw_emb, 16000, true, nullptr);
  } else {
    std::ifstream f(path, std::ios::binary);
    std::string buf((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    result = FunOfflineInferBuffer(asr, buf.data(), (int)buf.size(), RASR_NONE, nullptr, hw_emb, 16000, "pcm", true, nullptr);
  }
  if (!result) {
    fprintf(stderr, "inference failed\n");
    FunOfflineUninit(asr);
    return 1;
  }
  printf("text\t%s\nstamp\t%s\nsents\t%s\nseconds\t%.3f\nnumber\t%d\n", FunASRGetResult(result, 0), FunASRGetStamp(result),
         FunASRGetStampSents(result), FunASRGetRetSnippetTime(result), FunASRGetRetNumber(result));
  FunASRFreeResult(result);
  FunOfflineUninit(asr);
  return 0;
}
