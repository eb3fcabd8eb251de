// Shared helpers for libm3r_b200.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../../include/m3r_backends.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libm3r_b200 targets sm_100a (B200) only"
#endif

namespace m3r {

constexpr int kNumSMs = 148;  // B200

// thread-local error string behind m3r_last_error()
void set_error(const char *fmt, ...);
extern std::atomic<uint64_t> g_launch_count;

inline int cuda_fail(cudaError_t e, const char *what) {
  set_error("%s: %s", what, cudaGetErrorString(e));
  return M3R_ERR_CUDA;
}

#define M3R_CUDA_TRY(expr)                                  \
  do {                                                      \
    cudaError_t _e = (expr);                                \
    if (_e != cudaSuccess) return ::m3r::cuda_fail(_e, #expr); \
  } while (0)

// call right after a <<<>>> launch
#define M3R_LAUNCH_CHECK(name)                                   \
  do {                                                           \
    ::m3r::g_launch_count.fetch_add(1, std::memory_order_relaxed); \
    cudaError_t _e = cudaGetLastError();                         \
    if (_e != cudaSuccess) return ::m3r::cuda_fail(_e, name);    \
  } while (0)

#define M3R_REQUIRE(cond, ...)            \
  do {                                    \
    if (!(cond)) {                        \
      ::m3r::set_error(__VA_ARGS__);      \
      return M3R_ERR_INVALID_ARGUMENT;    \
    }                                     \
  } while (0)

inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace m3r
