// Library-level entry points: version, thread-local error text, launch counter.
#include <stdarg.h>

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
}  // namespace fb200

extern "C" __attribute__((visibility("default"))) int fb200_version(void) { return FB200_VERSION; }
extern "C" __attribute__((visibility("default"))) const char* fb200_last_error(void) { return fb200::g_err; }
extern "C" __attribute__((visibility("default"))) long long fb200_launch_count(int reset) {
  long long v = fb200::g_launches;
  if (reset) fb200::g_launches = 0;
  return v;
}
