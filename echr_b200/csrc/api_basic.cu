// C-ABI: library info, error string, fp32 -> bf16 plane conversion, and the stand-alone GEMM entry
// (used for the one-off projections and by the parity tests of the tensor-core kernel).
#include "../../include/echr_b200.h"
#include "engine.cuh"
#include "gemm_tc2.cuh"
#include "util.cuh"

namespace echr {
char* last_error_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}

}  // namespace echr

using namespace echr;

extern "C" {

const char* echr_last_error(void) { return last_error_buf(); }

int echr_abi_version(void) { return ECHR_B200_ABI_VERSION; }

int echr_device_check(int device) {
  cudaDeviceProp prop;
  ECHR_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return set_error(ECHR_ERR_NO_DEVICE, "device %d is sm_%d%d; this library is built for sm_100a only",
                     device, prop.major, prop.minor);
  return ECHR_OK;
}

int echr_split_bf16(const float* src, int64_t rows, int64_t cols, int64_t ld_src, void* hi, void* lo,
                    int64_t ld_dst, int64_t cols_dst, void* stream) {
  ECHR_REQUIRE(src && hi, "echr_split_bf16: null pointer");
  ECHR_REQUIRE(rows >= 0 && cols >= 0 && cols_dst >= cols && ld_dst >= cols_dst && ld_src >= cols,
               "echr_split_bf16: bad shape");
  ECHR_CUDA(launch_split(src, rows, cols, ld_src, (bf16*)hi, (bf16*)lo, ld_dst, cols_dst, false, (cudaStream_t)stream));
  return ECHR_OK;
}

int echr_split_f16(const float* src, int64_t rows, int64_t cols, int64_t ld_src, void* dst, int64_t ld_dst,
                   int64_t cols_dst, void* stream) {
  ECHR_REQUIRE(src && dst, "echr_split_f16: null pointer");
  ECHR_REQUIRE(rows >= 0 && cols >= 0 && cols_dst >= cols && ld_dst >= cols_dst && ld_src >= cols,
               "echr_split_f16: bad shape");
  ECHR_CUDA(launch_split(src, rows, cols, ld_src, (bf16*)dst, nullptr, ld_dst, cols_dst, true, (cudaStream_t)stream));
  return ECHR_OK;
}

int echr_attention_step(const float* q, const void* proj, const void* clip, int64_t ld_clip,
                        int clip_bf16, const float* alpha, const int32_t* row_start,
                        const int32_t* row_len, int64_t rows, float* out_f32, void* out_hi,
                        void* out_lo, int64_t ld_bf, int max_len, void* stream) {
  ECHR_REQUIRE(q && proj && clip && alpha && row_start && row_len, "echr_attention_step: null pointer");
  ECHR_REQUIRE(rows >= 0 && rows < (1ll << 31), "echr_attention_step: bad row count");
  ECHR_REQUIRE(ld_clip >= ECHR_VIDEO_DIM && (ld_clip * (clip_bf16 ? 2 : 4)) % 16 == 0,
               "echr_attention_step: clip row stride must be >= 500 and 16-byte aligned");
  ECHR_CUDA(launch_attention(q, proj, clip, ld_clip, clip_bf16, alpha, row_start, row_len, rows, out_f32,
                             (bf16*)out_hi, (bf16*)out_lo, ld_bf, (cudaStream_t)stream, max_len));
  return ECHR_OK;
}

int echr_attention_step_f16(const float* q_exp, const void* tab16, const float* proj_scale, const float* alpha,
                            const int32_t* row_start, const int32_t* row_len, int64_t rows, float* out_f32,
                            void* out16, int64_t ld_out, int mean_len, void* stream) {
  ECHR_REQUIRE(q_exp && tab16 && proj_scale && alpha && row_start && row_len, "echr_attention_step_f16: null pointer");
  ECHR_REQUIRE(rows >= 0 && rows < (1ll << 31), "echr_attention_step_f16: bad row count");
  ECHR_REQUIRE(((uintptr_t)tab16 & 15) == 0, "echr_attention_step_f16: the table must be 16-byte aligned");
  ECHR_REQUIRE(out16 == nullptr || (ld_out >= 512 && ld_out % 8 == 0 && ((uintptr_t)out16 & 15) == 0),
               "echr_attention_step_f16: out16 rows must be 16-byte aligned and >= 512 wide");
  AttF16Params P;
  P.q_exp = q_exp; P.tab = (const bf16*)tab16; P.proj_scale = proj_scale; P.alpha = alpha;
  P.row_start = row_start; P.row_len = row_len; P.rows = (int)rows;
  P.out = (bf16*)out16; P.ld_out = ld_out; P.out_f32 = out_f32;
  ECHR_CUDA(launch_attention_f16(P, mean_len, (cudaStream_t)stream));
  return ECHR_OK;
}

int echr_gemm_planes(const void* a, int64_t m, int64_t k, int64_t a_ld, int a_planes,
                     int64_t a_plane_stride, const void* w, int64_t n, int64_t w_ld, int w_planes,
                     int64_t w_plane_stride, const float* bias, int act, float* out_f32,
                     int64_t ld_f32, void* out_hi, void* out_lo, int64_t ld_bf, int block_n,
                     void* stream) {
  ECHR_REQUIRE(a && w, "echr_gemm_planes: null operand");
  ECHR_REQUIRE(m > 0 && n > 0 && k > 0 && k % GEMM_BLOCK_K == 0 && k / GEMM_BLOCK_K <= GEMM_MAX_KB,
               "echr_gemm_planes: K=%lld must be a multiple of 64 and <= %d", (long long)k,
               GEMM_BLOCK_K * GEMM_MAX_KB);
  ECHR_REQUIRE((a_planes == 1 || a_planes == 2) && (w_planes == 1 || w_planes == 2) &&
                   !(a_planes == 1 && w_planes == 2),
               "echr_gemm_planes: unsupported plane combination a=%d w=%d", a_planes, w_planes);
  ECHR_REQUIRE(block_n == 128 || block_n == 64 || block_n == 32 || block_n == 16 || block_n == 256, "echr_gemm_planes: block_n must be 16, 32, 64, 128 or 256 (persistent kernel)");
  ECHR_REQUIRE(m < (1ll << 31) && n < (1ll << 31), "echr_gemm_planes: dimension too large");
  GemmParams P;
  memset(&P, 0, sizeof(P));
  GemmProblem& pb = P.prob[0];
  const char* err = make_tmap_bf16(&pb.tmA, a, (uint64_t)k, (uint64_t)m, (uint64_t)a_planes,
                                   (uint64_t)a_ld, (uint64_t)a_plane_stride, GEMM_BLOCK_M);
  if (err) return set_error(ECHR_ERR_CUDA, "echr_gemm_planes: A: %s", err);
  err = make_tmap_bf16(&pb.tmW, w, (uint64_t)k, (uint64_t)n, (uint64_t)w_planes, (uint64_t)w_ld,
                       (uint64_t)w_plane_stride, (uint32_t)block_n);
  if (err) return set_error(ECHR_ERR_CUDA, "echr_gemm_planes: W: %s", err);
  pb.M = (int)m;
  pb.N = (int)n;
  pb.num_kb = (int)(k / GEMM_BLOCK_K);
  pb.act = act;
  for (int i = 0; i < pb.num_kb; ++i) pb.a_kb[i] = (uint16_t)i;
  pb.bias = bias;
  pb.out_f32 = out_f32;
  pb.ld_f32 = ld_f32;
  pb.out_hi = (__nv_bfloat16*)out_hi;
  pb.out_lo = (__nv_bfloat16*)out_lo;
  pb.ld_bf = ld_bf;
  if (block_n == 256) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    ECHR_CUDA(gemm2_launch(P, 1, a_planes, w_planes, EPI_STORE, 1, sms, (cudaStream_t)stream));
  } else
  ECHR_CUDA(gemm_launch(P, 1, (int)m, (int)n, a_planes, w_planes, block_n, EPI_STORE, 1,
                        (cudaStream_t)stream));
  count_launch();
  return ECHR_OK;
}

}  // extern "C"
