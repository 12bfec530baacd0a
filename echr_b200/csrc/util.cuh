// Error plumbing shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <utility>

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

// Programmatic dependent launch: the kernel may start (prologue only) while its stream predecessor
// is still running; it calls pdl_wait() before touching anything the predecessor writes.
// ECHR_NO_PDL=1 in the environment falls back to plain stream order.
// ECHR_PDL_MASK selects which step kernels use it (bit 0 h2att, 1 attention, 2 gates, 3 logits,
// 4 finalize); kernels with large shared memory that become resident early can starve a
// concurrent graph branch: measured on cfg2 every non-zero mask was slower than plain stream order
// inside the fork/join decode graph (profiles/r01_pdl_sweep.md), so the default is off.
enum { PDL_H2ATT = 1, PDL_ATT = 2, PDL_GATES = 4, PDL_LOGIT = 8, PDL_FINALIZE = 16, PDL_DEFAULT_MASK = 0 };
inline int pdl_mask() {
  static int mask = -1;
  if (mask < 0) {
    const char* e = getenv("ECHR_PDL_MASK");
    mask = e ? atoi(e) : PDL_DEFAULT_MASK;
    const char* off = getenv("ECHR_NO_PDL");
    if (off != nullptr && off[0] == '1') mask = 0;
  }
  return mask;
}
inline bool& pdl_next() {  // whether the next launch_pdl() on this thread may use the attribute
  static thread_local bool on = false;
  return on;
}
struct PdlScope {
  bool prev;
  explicit PdlScope(int bit) : prev(pdl_next()) { pdl_next() = (pdl_mask() & bit) != 0; }
  ~PdlScope() { pdl_next() = prev; }
};
inline bool pdl_enabled() { return pdl_next(); }
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

}  // namespace echr
