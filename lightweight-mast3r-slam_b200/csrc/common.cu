// Error string, launch counter and version for libm3r_b200.so.
#include "m3r_common.cuh"

#include <cstring>

namespace m3r {

static thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launch_count{0};

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

}  // namespace m3r

extern "C" int m3r_abi_version(void) { return M3R_ABI_VERSION; }
extern "C" const char *m3r_last_error(void) { return m3r::g_err; }
extern "C" uint64_t m3r_launch_count(void) {
  return m3r::g_launch_count.load(std::memory_order_relaxed);
}
