// Library-level entry points: version, thread-local error text, launch counter.
#include <stdarg.h>
#include <stdlib.h>

#include <vector>

#include "common.cuh"

namespace fb200 {
static thread_local char g_err[512] = "";
static thread_local long long g_launches = 0;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches += n; }

bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("FB200_PDL");
    return e == nullptr || atoi(e) != 0;
  }();
  return on;
}

// ---- optional per-entry-point device timing (CUDA events on the caller's stream) ----------
struct ProfRec {
  cudaEvent_t beg, end;
  int kind;
  double work;
};
static thread_local bool g_prof_on = false;
static thread_local std::vector<ProfRec>* g_prof = nullptr;

ProfScope::ProfScope(int kind, void* stream, double work) : active_(g_prof_on), stream_(stream) {
  if (!active_) return;
  if (g_prof == nullptr) g_prof = new std::vector<ProfRec>();
  ProfRec r;
  r.kind = kind;
  r.work = work;
  if (cudaEventCreate(&r.beg) != cudaSuccess || cudaEventCreate(&r.end) != cudaSuccess) {
    active_ = false;
    return;
  }
  cudaEventRecord(r.beg, (cudaStream_t)stream);
  g_prof->push_back(r);
}
ProfScope::~ProfScope() {
  if (active_) cudaEventRecord(g_prof->back().end, (cudaStream_t)stream_);
}
}  // namespace fb200

extern "C" __attribute__((visibility("default"))) int fb200_profile_enable(int on) {
  fb200::g_prof_on = on != 0;
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_profile_read(int kind, double* ms, long long* calls,
                                                                          double* work, int reset) {
  double tot = 0.0, wk = 0.0;
  long long n = 0;
  if (fb200::g_prof != nullptr) {
    for (auto& r : *fb200::g_prof) {
      if (r.kind != kind) continue;
      if (cudaEventSynchronize(r.end) != cudaSuccess) {
        fb200::set_error("profile_read: cudaEventSynchronize failed");
        return FB200_ECUDA;
      }
      float t = 0.f;
      cudaEventElapsedTime(&t, r.beg, r.end);
      tot += t;
      wk += r.work;
      ++n;
    }
    if (reset) {
      auto& v = *fb200::g_prof;
      size_t keep = 0;
      for (size_t i = 0; i < v.size(); ++i) {
        if (v[i].kind == kind) {
          cudaEventDestroy(v[i].beg);
          cudaEventDestroy(v[i].end);
        } else {
          v[keep++] = v[i];
        }
      }
      v.resize(keep);
    }
  }
  if (ms) *ms = tot;
  if (calls) *calls = n;
  if (work) *work = wk;
  return FB200_OK;
}

extern "C" __attribute__((visibility("default"))) int fb200_version(void) { return FB200_VERSION; }
extern "C" __attribute__((visibility("default"))) const char* fb200_last_error(void) { return fb200::g_err; }
extern "C" __attribute__((visibility("default"))) long long fb200_launch_count(int reset) {
  long long v = fb200::g_launches;
  if (reset) fb200::g_launches = 0;
  return v;
}
