/* echr_b200 — C-ABI of the B200-native (sm_100a) ECHR caption-generation hot path.
 *
 * The reference (ttengwang/ECHR) is pure Python/PyTorch and has no FFI of its own; the entry
 * points below are the ones a binding for its hot path would need.  Each one names the reference
 * code it replaces (paths relative to the reference repository root).
 *
 * Conventions
 *   - every function returns 0 on success or a negative code; echr_last_error() gives the text
 *     (thread-local).  No function allocates behind the caller's back except the *_create calls.
 *   - all pointers are DEVICE pointers unless the name says host; sizes/strides are in elements.
 *   - `stream` is a cudaStream_t passed as void*; work is stream-ordered, nothing synchronises
 *     unless stated.
 *   - there is no CPU fallback: on a non-sm_100 device every entry fails with ECHR_ERR_NO_DEVICE.
 */
#ifndef ECHR_B200_H_
#define ECHR_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ECHR_B200_ABI_VERSION 1

#if defined(__GNUC__)
#define ECHR_API __attribute__((visibility("default")))
#else
#define ECHR_API
#endif

/* ---- library ---------------------------------------------------------------------------- */
ECHR_API const char* echr_last_error(void);
ECHR_API int echr_abi_version(void);
/* 0 if `device` is an sm_100 GPU this build can run on. */
ECHR_API int echr_device_check(int device);

/* ---- building blocks ---------------------------------------------------------------------- */
/* fp32 [rows, cols] (row stride ld_src) -> bf16 planes hi (and lo = bf16(x - hi), may be NULL),
 * zero-padded to cols_dst columns, row stride ld_dst.  Used to stage tensor-core operands. */
ECHR_API int echr_split_bf16(const float* src, int64_t rows, int64_t cols, int64_t ld_src, void* hi, void* lo,
                    int64_t ld_dst, int64_t cols_dst, void* stream);

/* out[m, n] = act( A[m, k] * W[n, k]^T + bias[n] ) on the tcgen05 tensor cores.
 * Replaces every nn.Linear on the path (models/MA_attention_8_NEW.py:29,93-96;
 * models/OldModel_NEW.py:51,372-374).  A and W are bf16 plane stacks ([planes][rows][ld]):
 * planes=1 is the plain bf16 operand, planes=2 adds the low-order plane (fp32-grade product).
 * k must be a multiple of 64 (pad with zeros), k <= 2048.  act: 0 none, 1 tanh, 2 tanh.approx.
 * Outputs: out_f32 (may be NULL) and/or bf16 planes out_hi/out_lo (may be NULL). */
ECHR_API int echr_gemm_planes(const void* a, int64_t m, int64_t k, int64_t a_ld, int a_planes,
                     int64_t a_plane_stride, const void* w, int64_t n, int64_t w_ld, int w_planes,
                     int64_t w_plane_stride, const float* bias, int act, float* out_f32,
                     int64_t ld_f32, void* out_hi, void* out_lo, int64_t ld_bf, int block_n,
                     void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ECHR_B200_H_ */
