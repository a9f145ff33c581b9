// Error reporting, version and launch accounting for libzerotig.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include "common.cuh"

namespace zt {
static thread_local char g_err[512] = "";
std::atomic<int64_t> g_launches{0};
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
bool pdl_enabled() {
  static const bool on = getenv("ZT_NO_PDL") == nullptr;
  return on;
}
}  // namespace zt

extern "C" const char* zt_last_error(void) { return zt::g_err; }
extern "C" int zt_version(void) { return 100; }
extern "C" int64_t zt_launch_count(void) { return zt::g_launches.load(); }
/* kernels replayed from a captured CUDA graph are counted by the host (they do not pass through check_launch) */
extern "C" int zt_add_launches(int64_t n) {
  zt::g_launches.fetch_add(n, std::memory_order_relaxed);
  return ZT_OK;
}
