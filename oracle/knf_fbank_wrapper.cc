// TEST INFRASTRUCTURE (checker only -- see oracle/__init__.py).
// C entry point around the reference's OWN vendored fbank (kaldi-native-fbank, compiled from where it lies under
// /root/reference by oracle/Makefile; nothing of it is copied into this repository).  The call sequence is the one
// of Paraformer::FbankKaldi / Paraformer::InitAsr in the reference's C++ runtime
// (runtime/onnxruntime/src/paraformer.cpp:20-31 options, :298-311 feeding): dither 0, energy_floor 0, samples x 32768,
// OnlineFbank::AcceptWaveform, frames read back one by one.  Used by tests/test_oracle_knf.py as a SECOND, independent
// pin of the fbank stage next to torchaudio.compliance.kaldi.
#include <cstdint>
#include <string>
#include <vector>

#include "kaldi-native-fbank/csrc/online-feature.h"

extern "C" int knf_fbank(const float* wave, int num_samples, float sample_rate, int n_mels, float frame_length_ms,
                         float frame_shift_ms, const char* window_type, float* out, int max_frames) {
  knf::FbankOptions opts;
  opts.frame_opts.dither = 0;
  opts.mel_opts.num_bins = n_mels;
  opts.frame_opts.samp_freq = sample_rate;
  opts.frame_opts.window_type = window_type;
  opts.frame_opts.frame_shift_ms = frame_shift_ms;
  opts.frame_opts.frame_length_ms = frame_length_ms;
  opts.energy_floor = 0;
  opts.mel_opts.debug_mel = false;
  knf::OnlineFbank fbank(opts);
  std::vector<float> buf(num_samples);
  for (int i = 0; i != num_samples; ++i) buf[i] = wave[i] * 32768;
  fbank.AcceptWaveform(sample_rate, buf.data(), (int32_t)buf.size());
  const int frames = fbank.NumFramesReady();
  if (out == nullptr) return frames;
  const int n = frames < max_frames ? frames : max_frames;
  for (int i = 0; i != n; ++i) {
    const float* f = fbank.GetFrame(i);
    for (int m = 0; m != n_mels; ++m) out[(size_t)i * n_mels + m] = f[m];
  }
  return frames;
}
