// Error plumbing shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>

namespace echr {

// codes returned through the C-ABI (0 = ok)
enum : int {
  ECHR_OK = 0,
  ECHR_ERR_INVALID = -1,   // bad argument / unsupported configuration
  ECHR_ERR_CUDA = -2,      // a CUDA runtime / driver call failed
  ECHR_ERR_NO_DEVICE = -3, // no sm_100 device
  ECHR_ERR_WORKSPACE = -4  // workspace too small / not initialised
};

char* last_error_buf();  // thread-local, 512 bytes

inline int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(last_error_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}

#define ECHR_CUDA(expr)                                                                     \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess)                                                                  \
      return ::echr::set_error(::echr::ECHR_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,       \
                               cudaGetErrorString(_e), __FILE__, __LINE__);                 \
  } while (0)

#define ECHR_REQUIRE(cond, ...)                                                             \
  do {                                                                                      \
    if (!(cond)) return ::echr::set_error(::echr::ECHR_ERR_INVALID, __VA_ARGS__);           \
  } while (0)

}  // namespace echr
