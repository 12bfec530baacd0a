// Host side of the tcgen05 GEMM: tensor-map encoding (driver entry point fetched at run time so
// the library has no link-time libcuda dependency) and the template dispatch table.
#include "gemm_tc.cuh"
#include "gemm_tc2.cuh"
#include "util.cuh"

#include <mutex>

namespace echr {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

const char* make_tmap_bf16(CUtensorMap* out, const void* ptr, uint64_t cols, uint64_t rows,
                           uint64_t planes, uint64_t ld_elems, uint64_t plane_stride_elems,
                           uint32_t box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return "cuTensorMapEncodeTiled entry point not available";
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0) return "tensor map base not 16-byte aligned";
  if ((ld_elems * 2) % 16 != 0 || (plane_stride_elems * 2) % 16 != 0)
    return "tensor map strides must be multiples of 16 bytes";
  if (box_rows == 0 || box_rows > 256) return "tensor map box rows out of range";
  cuuint64_t gdim[3] = {cols, rows, planes};
  cuuint64_t gstr[2] = {ld_elems * 2, plane_stride_elems * 2};
  cuuint32_t box[3] = {static_cast<cuuint32_t>(GEMM_BLOCK_K), box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), gdim, gstr, box,
                  estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return "cuTensorMapEncodeTiled failed";
  return nullptr;
}

template <int A, int W, int BN, int EPI, int TOPK>
static cudaError_t launch_one(const GemmParams& P, dim3 grid, cudaStream_t stream) {
  using Cfg = GemmCfg<A, W, BN>;
  auto kern = gemm_tc_kernel<A, W, BN, EPI, TOPK>;
  static int attr[MAX_DEVICES] = {};  // per instantiation and device
  if (cudaError_t e = set_max_dynamic_smem(kern, Cfg::SMEM_BYTES, attr); e != cudaSuccess) return e;
  return launch_pdl(kern, grid, dim3(GEMM_THREADS), Cfg::SMEM_BYTES, stream, P);
}

template <int A, int W>
static cudaError_t launch_aw(const GemmParams& P, dim3 grid, int block_n, int epi, int topk,
                             cudaStream_t stream) {
  if (epi == EPI_STORE && block_n == 128) return launch_one<A, W, 128, EPI_STORE, 1>(P, grid, stream);
  if (epi == EPI_STORE && block_n == 64) return launch_one<A, W, 64, EPI_STORE, 1>(P, grid, stream);
  if (epi == EPI_STORE && block_n == 32) return launch_one<A, W, 32, EPI_STORE, 1>(P, grid, stream);
  if (epi == EPI_STORE && block_n == 16) return launch_one<A, W, 16, EPI_STORE, 1>(P, grid, stream);
  if (epi == EPI_LSTM && block_n == 128) return launch_one<A, W, 128, EPI_LSTM, 1>(P, grid, stream);
  if (epi == EPI_LOGITS && block_n == 128 && topk == 1)
    return launch_one<A, W, 128, EPI_LOGITS, 1>(P, grid, stream);
  if (epi == EPI_LOGITS && block_n == 128 && topk == 4)
    return launch_one<A, W, 128, EPI_LOGITS, 4>(P, grid, stream);
  return cudaErrorInvalidValue;
}

cudaError_t gemm_launch(const GemmParams& P, int groups, int max_m, int max_n, int a_terms,
                        int w_terms, int block_n, int epi, int topk, cudaStream_t stream) {
  if (groups < 1 || groups > GEMM_MAX_GROUPS || max_m <= 0 || max_n <= 0)
    return cudaErrorInvalidValue;
  dim3 grid((max_n + block_n - 1) / block_n, (max_m + GEMM_BLOCK_M - 1) / GEMM_BLOCK_M, groups);
  if (a_terms == 1 && w_terms == 1) return launch_aw<1, 1>(P, grid, block_n, epi, topk, stream);
  if (a_terms == 2 && w_terms == 1) return launch_aw<2, 1>(P, grid, block_n, epi, topk, stream);
  if (a_terms == 2 && w_terms == 2) return launch_aw<2, 2>(P, grid, block_n, epi, topk, stream);
  return cudaErrorInvalidValue;
}

// ---------------------------------------------------------------- cluster split-K variant
template <int A, int W, int BN, int EPI, int TOPK>
static cudaError_t launch_split_one(const GemmParams& P, dim3 grid, cudaStream_t stream) {
  using Cfg = GemmCfg<A, W, BN>;
  auto kern = gemm_tc_kernel<A, W, BN, EPI, TOPK, 2>;
  static int attr[MAX_DEVICES] = {};  // per instantiation and device
  if (cudaError_t e = set_max_dynamic_smem(kern, Cfg::SMEM_BYTES, attr); e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, kern, P);
}

template <int A, int W>
static cudaError_t launch_split_aw(const GemmParams& P, dim3 grid, int block_n, int epi, int topk, cudaStream_t stream) {
  if (epi == EPI_LSTM && block_n == 128) return launch_split_one<A, W, 128, EPI_LSTM, 1>(P, grid, stream);
  if (epi == EPI_LOGITS && block_n == 256 && topk == 1) return launch_split_one<A, W, 256, EPI_LOGITS, 1>(P, grid, stream);
  if (epi == EPI_LOGITS && block_n == 256 && topk == 4) return launch_split_one<A, W, 256, EPI_LOGITS, 4>(P, grid, stream);
  return cudaErrorInvalidValue;
}

cudaError_t gemm_launch_split(const GemmParams& P, int groups, int max_m, int grid_x, int a_terms, int w_terms,
                              int block_n, int epi, int topk, cudaStream_t stream) {
  if (groups < 1 || groups > GEMM_MAX_GROUPS || max_m <= 0 || grid_x <= 0) return cudaErrorInvalidValue;
  dim3 grid((grid_x + 1) & ~1, (max_m + GEMM_BLOCK_M - 1) / GEMM_BLOCK_M, groups);
  if (a_terms == 1 && w_terms == 1) return launch_split_aw<1, 1>(P, grid, block_n, epi, topk, stream);
  if (a_terms == 2 && w_terms == 1) return launch_split_aw<2, 1>(P, grid, block_n, epi, topk, stream);
  if (a_terms == 2 && w_terms == 2) return launch_split_aw<2, 2>(P, grid, block_n, epi, topk, stream);
  return cudaErrorInvalidValue;
}

// ---------------------------------------------------------------- persistent 128 x 256 kernel
template <int A, int W, int EPI, int TOPK>
static cudaError_t launch_two(const GemmParams& P, int groups, int total_tiles, int num_sms, cudaStream_t stream) {
  using Cfg = Gemm2Cfg<A, W, gemm2_lstm_staged(A, W, EPI)>;
  const int grid = total_tiles < num_sms ? total_tiles : num_sms;
  auto kern = gemm_tc2_kernel<A, W, EPI, TOPK>;
  static int attr[MAX_DEVICES] = {};  // per instantiation and device
  if (cudaError_t e = set_max_dynamic_smem(kern, Cfg::SMEM_BYTES, attr); e != cudaSuccess) return e;
  return launch_pdl(kern, dim3(grid), dim3(GEMM_THREADS), Cfg::SMEM_BYTES, stream, P, groups, total_tiles);
}

template <int A, int W>
static cudaError_t launch_two_aw(const GemmParams& P, int groups, int total, int epi, int topk, int num_sms,
                                 cudaStream_t stream) {
  if (epi == EPI_STORE) return launch_two<A, W, EPI_STORE, 1>(P, groups, total, num_sms, stream);
  if (epi == EPI_LSTM) return launch_two<A, W, EPI_LSTM, 1>(P, groups, total, num_sms, stream);
  if (epi == EPI_LOGITS && topk == 1) return launch_two<A, W, EPI_LOGITS, 1>(P, groups, total, num_sms, stream);
  if (epi == EPI_LOGITS && topk == 4) return launch_two<A, W, EPI_LOGITS, 4>(P, groups, total, num_sms, stream);
  return cudaErrorInvalidValue;
}

cudaError_t gemm2_launch(const GemmParams& P, int groups, int a_terms, int w_terms, int epi, int topk,
                         int num_sms, cudaStream_t stream) {
  if (groups < 1 || groups > GEMM_MAX_GROUPS) return cudaErrorInvalidValue;
  int total = 0;
  for (int g = 0; g < groups; ++g)
    total += ((P.prob[g].M + GEMM_BLOCK_M - 1) / GEMM_BLOCK_M) * ((P.prob[g].N + GEMM2_BLOCK_N - 1) / GEMM2_BLOCK_N);
  if (total <= 0) return cudaSuccess;
  if (a_terms == 1 && w_terms == 1) return launch_two_aw<1, 1>(P, groups, total, epi, topk, num_sms, stream);
  if (a_terms == 2 && w_terms == 1) return launch_two_aw<2, 1>(P, groups, total, epi, topk, num_sms, stream);
  if (a_terms == 2 && w_terms == 2) return launch_two_aw<2, 2>(P, groups, total, epi, topk, num_sms, stream);
  return cudaErrorInvalidValue;
}

}  // namespace echr
