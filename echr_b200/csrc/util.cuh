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

// Experiment knobs (ECHR_TRACE, ECHR_KSPLIT, ECHR_PDL_MASK, ...: the variables named next to each knob() call)
// exist only in a debug build, `make DEBUG_KNOBS=1`: the shipped library reads no environment variable, every
// schedule choice is a function of the batch shape alone.
#ifdef ECHR_DEBUG_KNOBS
inline const char* knob(const char* name) { return getenv(name); }
#else
inline const char* knob(const char*) { return nullptr; }
#endif

// ---- per-device launch state.  cudaFuncSetAttribute and the SM count are per DEVICE; one process may drive
// several GPUs (Engine(device=...)), so "already set" is tracked per device ordinal, not per process.
constexpr int MAX_DEVICES = 64;

// Opt a kernel in to `bytes` of dynamic shared memory on the current device (no-op once done there).
// `state` is a zero-initialised int[MAX_DEVICES] owned by the caller, one per kernel instantiation.
template <typename K>
inline cudaError_t set_max_dynamic_smem(K kern, int bytes, int* state) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= MAX_DEVICES) return cudaErrorInvalidDevice;
  if (state[dev] >= bytes) return cudaSuccess;
  e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) state[dev] = bytes;
  return e;
}

inline int num_sms() {  // of the current device
  static int n[MAX_DEVICES] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) return 148;
  if (n[dev] == 0) {
    int v = 0;
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    n[dev] = v > 0 ? v : 148;
  }
  return n[dev];
}

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
    const char* e = knob("ECHR_PDL_MASK");
    mask = e ? atoi(e) : PDL_DEFAULT_MASK;
    const char* off = knob("ECHR_NO_PDL");
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
