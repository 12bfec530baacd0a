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
// ECHR_PDL_MASK selects which step kernels are launched as programmatic dependents (bit 0 h2att, 1 attention,
// 2 gates, 3 logits, 4 finalize).  What matters is WHEN the primary releases them: released right after the
// primary's prologue (ECHR_PDL_LATE=0) the dependents' CTAs sit on SMs (and ~200 KB of shared memory each) for
// the whole primary and starve the concurrent graph branch — every mask is slower than plain stream order
// (cfg2: 186k vs 232k events/s).  Released when the primary's accumulators are ready (the default), only the
// launch latency and the prologue of the dependent overlap the primary's epilogue: 232k -> 254k events/s.
// Default mask 7 = the recurrent chain gates -> h2att -> attention -> gates (one stream).  The logit GEMM
// follows an event from another stream, where the attribute has no primary to hang on; finalize stays a plain
// launch: as a programmatic dependent of the logits it sat between them and the event the next gate GEMM waits
// for, and that GEMM then lost its own early launch (cfg2 with arg-max keys: 250k with it, 265k without).
enum { PDL_H2ATT = 1, PDL_ATT = 2, PDL_GATES = 4, PDL_LOGIT = 8, PDL_FINALIZE = 16, PDL_DEFAULT_MASK = 7 };
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
inline bool& pdl_regime() {  // set by the decode enqueue: programmatic launches pay off in the latency regime only
  static thread_local bool on = true;
  return on;
}
struct PdlScope {
  bool prev;
  explicit PdlScope(int bit) : prev(pdl_next()) { pdl_next() = pdl_regime() && (pdl_mask() & bit) != 0; }
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
