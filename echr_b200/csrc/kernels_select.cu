// Proposal selection between the proposal module and the caption path, one CTA per video:
//   select_topn_kernel   eval_utils.py:263-292  gettop1000      radix select of the topN-th score + ordered compaction
//   select_nms_kernel    eval_utils.py:295-331  gettop1000_nms  greedy 1-D NMS over the (n,k) proposal grid
//   rerank_nms_kernel    eval_utils.py:230-260  gettopN_nms     NMS that keeps the best-sentence member of each cluster
// The reference walks the [T,K] score matrix in Python double loops on the host after a D->H copy; here the
// matrix never leaves HBM and only the selected (ind, start, end, score) records are produced.
// Integer / comparison work: results are bit-exact with the numpy code (overlaps are evaluated in fp64 with the
// same operation order; IEEE division is correctly rounded on both sides).
#include "engine.cuh"
#include "util.cuh"

namespace echr {

constexpr int SEL_THREADS = 1024;

// order-preserving map float -> uint32 (larger float = larger key); no real float maps to 0
__device__ __forceinline__ uint32_t order_key(float x) {
  const uint32_t u = __float_as_uint(x);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_value(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k);
}

__device__ __forceinline__ unsigned long long block_max_u64(unsigned long long v, unsigned long long* smem) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long w = __shfl_xor_sync(0xffffffffu, v, o);
    v = w > v ? w : v;
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();  // smem reuse across calls
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  v = smem[lane];  // SEL_THREADS / 32 == 32 warps
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long w = __shfl_xor_sync(0xffffffffu, v, o);
    v = w > v ? w : v;
  }
  return v;
}

// ---------------------------------------------------------------------------------------------
// gettop1000: p = scores*mask; thr = topn-th largest p over ALL T*K cells (the smallest when there are fewer);
// emit, in (n,k) row-major order, the cells with n >= k and p >= max(thr, val_thres).
__global__ void __launch_bounds__(SEL_THREADS)
select_topn_kernel(const float* __restrict__ scores, const float* __restrict__ mask,
                   const int32_t* __restrict__ clip_off, int K, int topn, float val_thres,
                   int32_t* __restrict__ out_ind, int32_t* __restrict__ out_start, int32_t* __restrict__ out_end,
                   float* __restrict__ out_conf, int32_t* __restrict__ out_count, int cap) {
  __shared__ unsigned hist[256];
  __shared__ unsigned s_prefix, s_rank;
  __shared__ int s_warp[32];
  __shared__ int s_base;
  const int v = blockIdx.x;
  const long long row0 = clip_off[v];
  const int T = clip_off[v + 1] - clip_off[v];
  const long long N = (long long)T * K;
  if (N <= 0) {
    if (threadIdx.x == 0) out_count[v] = 0;
    return;
  }
  const float* sc = scores + row0 * K;
  const float* mk = mask ? mask + row0 * K : nullptr;
  auto cell = [&](long long i) { return mk ? sc[i] * mk[i] : sc[i]; };
  if (threadIdx.x == 0) {
    s_prefix = 0;
    s_rank = (unsigned)(N < topn ? N : topn);  // rank from the top, 1-based
  }
  for (int shift = 24; shift >= 0; shift -= 8) {
    if (threadIdx.x < 256) hist[threadIdx.x] = 0;
    __syncthreads();
    const unsigned prefix = s_prefix;
    const unsigned high = shift == 24 ? 0u : (0xFFFFFFFFu << (shift + 8));
    for (long long i = threadIdx.x; i < N; i += SEL_THREADS) {
      const uint32_t k = order_key(cell(i));
      if ((k & high) == prefix) atomicAdd(&hist[(k >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned r = s_rank;
      int b = 255;
      for (; b > 0; --b) {
        if (hist[b] >= r) break;
        r -= hist[b];
      }
      s_prefix = prefix | ((unsigned)b << shift);
      s_rank = r;
    }
    __syncthreads();
  }
  const float thr = fmaxf(key_value(s_prefix), val_thres);
  // ordered compaction: 4 consecutive cells per thread, chunks of 4096 cells
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  int32_t* oi = out_ind + (long long)v * cap;
  int32_t* os = out_start + (long long)v * cap;
  int32_t* oe = out_end + (long long)v * cap;
  float* oc = out_conf + (long long)v * cap;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (long long c0 = 0; c0 < N; c0 += 4LL * SEL_THREADS) {
    const long long i0 = c0 + 4LL * threadIdx.x;
    float p[4];
    bool ok[4];
    int mine = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long i = i0 + j;
      ok[j] = false;
      if (i < N) {
        p[j] = cell(i);
        const int n = (int)(i / K), k = (int)(i - (long long)n * K);
        ok[j] = (n >= k) && (p[j] >= thr);
      }
      mine += ok[j] ? 1 : 0;
    }
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int wsum = s_warp[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, wsum, o);
      if (lane >= o) wsum += t;
    }
    const int warp_excl = __shfl_sync(0xffffffffu, wsum, warp) - s_warp[warp];
    const int total = __shfl_sync(0xffffffffu, wsum, 31);
    int pos = s_base + warp_excl + incl - mine;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (ok[j]) {
        if (pos < cap) {
          const long long i = i0 + j;
          const int n = (int)(i / K), k = (int)(i - (long long)n * K);
          oi[pos] = n;
          os[pos] = n - k;
          oe[pos] = n + 1;
          oc[pos] = p[j];
        }
        ++pos;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) s_base += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) out_count[v] = s_base;
}

// ---------------------------------------------------------------------------------------------
// gettop1000_nms: proposals p <-> (n,k), n in [0,T), k in [0, min(n,K)), in that enumeration order:
// [start,end) = [n-k, n+1), score = scores[n,k].  Greedy: take the best remaining, drop every remaining proposal
// whose overlap with it exceeds `overlap`.  Ties in the score go to the larger enumeration index (what a stable
// ascending sort followed by ind[-1] gives; numpy's default quicksort leaves equal scores unordered).
__device__ __forceinline__ long long prop_base(int n, int K) {
  return n <= K + 1 ? (long long)n * (n - 1) / 2 : (long long)K * (K + 1) / 2 + (long long)(n - 1 - K) * K;
}

__global__ void __launch_bounds__(SEL_THREADS)
select_nms_kernel(const float* __restrict__ scores, const int32_t* __restrict__ clip_off, int K, double overlap,
                  int topn, uint32_t* __restrict__ ws_key, uint32_t* __restrict__ ws_nk,
                  const long long* __restrict__ ws_off, int32_t* __restrict__ out_ind,
                  int32_t* __restrict__ out_start, int32_t* __restrict__ out_end, float* __restrict__ out_score,
                  int32_t* __restrict__ out_pick, int32_t* __restrict__ out_count, int cap) {
  __shared__ unsigned long long s_red[32];
  const int v = blockIdx.x;
  const long long row0 = clip_off[v];
  const int T = clip_off[v + 1] - clip_off[v];
  const long long P = T > 0 ? prop_base(T, K) : 0;
  uint32_t* key = ws_key + ws_off[v];
  uint32_t* nk = ws_nk + ws_off[v];
  // enumerate: cell (n,k) -> p
  for (long long i = threadIdx.x; i < (long long)T * K; i += SEL_THREADS) {
    const int n = (int)(i / K), k = (int)(i - (long long)n * K);
    if (k < min(n, K)) {
      const long long p = prop_base(n, K) + k;
      key[p] = order_key(scores[(row0 + n) * K + k]);
      nk[p] = ((uint32_t)n << 16) | (uint32_t)k;
    }
  }
  __syncthreads();
  unsigned long long local = 0;
  for (long long p = threadIdx.x; p < P; p += SEL_THREADS) {
    const unsigned long long c = ((unsigned long long)key[p] << 32) | (unsigned long long)p;
    local = c > local ? c : local;
  }
  unsigned long long best = block_max_u64(local, s_red);
  int npick = 0;
  int32_t* oi = out_ind + (long long)v * cap;
  int32_t* os = out_start + (long long)v * cap;
  int32_t* oe = out_end + (long long)v * cap;
  int32_t* op = out_pick + (long long)v * cap;
  float* oc = out_score + (long long)v * cap;
  while ((best >> 32) != 0 && npick < topn && npick < cap) {
    const long long i = (long long)(best & 0xFFFFFFFFull);
    const uint32_t nki = nk[i];
    const int ni = (int)(nki >> 16), ki = (int)(nki & 0xFFFFu);
    const double t1i = ni - ki, t2i = ni + 1;
    const double area_i = t2i - t1i + 1.0;
    if (threadIdx.x == 0) {
      oi[npick] = ni;
      os[npick] = ni - ki;
      oe[npick] = ni + 1;
      oc[npick] = key_value((uint32_t)(best >> 32));
      op[npick] = (int32_t)i;
    }
    ++npick;
    local = 0;
    for (long long p = threadIdx.x; p < P; p += SEL_THREADS) {
      const uint32_t kp = key[p];
      if (kp == 0) continue;
      if (p == i) {
        key[p] = 0;
        continue;
      }
      const uint32_t q = nk[p];
      const int n = (int)(q >> 16), k = (int)(q & 0xFFFFu);
      const double t1 = n - k, t2 = n + 1;
      const double tt1 = fmax(t1i, t1), tt2 = fmin(t2i, t2);
      const double wh = fmax(0.0, tt2 - tt1 + 1.0);
      const double o = wh / (area_i + (t2 - t1 + 1.0) - wh);
      if (o <= overlap) {
        const unsigned long long c = ((unsigned long long)kp << 32) | (unsigned long long)p;
        local = c > local ? c : local;
      } else {
        key[p] = 0;
      }
    }
    best = block_max_u64(local, s_red);
  }
  if (threadIdx.x == 0) out_count[v] = npick;
}

// ---------------------------------------------------------------------------------------------
// gettopN_nms: while proposals remain: i = the best remaining by prop_score; cluster = remaining with
// overlap(i, .) >= nms_overlap; pick the cluster member with the largest sent_score (first in ascending
// prop_score order on ties); remaining = those with overlap(i, .) <= nms_overlap.  fp64 throughout like numpy.
struct RerankCand {
  double a, b;  // primary key (larger wins), secondary key (SMALLER wins when `second_small`)
  int idx;
};
__device__ __forceinline__ bool better_first(const RerankCand& x, const RerankCand& y) {
  // best remaining by prop_score: larger a, then larger index (stable ascending sort, last element)
  if (x.idx < 0) return false;
  if (y.idx < 0) return true;
  if (x.a != y.a) return x.a > y.a;
  return x.idx > y.idx;
}
__device__ __forceinline__ bool better_pick(const RerankCand& x, const RerankCand& y) {
  // largest sent_score; on ties the first in ascending (prop_score, index) order
  if (x.idx < 0) return false;
  if (y.idx < 0) return true;
  if (x.a != y.a) return x.a > y.a;
  if (x.b != y.b) return x.b < y.b;
  return x.idx < y.idx;
}
template <bool PICK>
__device__ __forceinline__ RerankCand block_best(RerankCand c, RerankCand* smem) {
  __syncthreads();
  smem[threadIdx.x] = c;
  __syncthreads();
  for (int s = SEL_THREADS / 2; s > 0; s >>= 1) {
    if (threadIdx.x < s) {
      const RerankCand o = smem[threadIdx.x + s];
      const RerankCand m = smem[threadIdx.x];
      if (PICK ? better_pick(o, m) : better_first(o, m)) smem[threadIdx.x] = o;
    }
    __syncthreads();
  }
  return smem[0];
}

__global__ void __launch_bounds__(SEL_THREADS)
rerank_nms_kernel(const double* __restrict__ t1, const double* __restrict__ t2, const double* __restrict__ prop,
                  const double* __restrict__ sent, const int32_t* __restrict__ off, double nms_overlap, int topn,
                  uint8_t* __restrict__ ws_alive, int32_t* __restrict__ out_pick, int32_t* __restrict__ out_count,
                  int cap) {
  extern __shared__ __align__(16) unsigned char raw[];
  RerankCand* smem = reinterpret_cast<RerankCand*>(raw);
  const int v = blockIdx.x;
  const int o0 = off[v], n = off[v + 1] - off[v];
  const double* a1 = t1 + o0;
  const double* a2 = t2 + o0;
  const double* ps = prop + o0;
  const double* ss = sent + o0;
  uint8_t* alive = ws_alive + o0;
  for (int j = threadIdx.x; j < n; j += SEL_THREADS) alive[j] = 1;
  int npick = 0;
  int32_t* op = out_pick + (long long)v * cap;
  while (npick < topn && npick < cap) {
    RerankCand c = {0.0, 0.0, -1};
    for (int j = threadIdx.x; j < n; j += SEL_THREADS)
      if (alive[j]) {
        const RerankCand d = {ps[j], 0.0, j};
        if (better_first(d, c)) c = d;
      }
    const RerankCand first = block_best<false>(c, smem);
    if (first.idx < 0) break;
    const int i = first.idx;
    const double t1i = a1[i], t2i = a2[i];
    const double area_i = t2i - t1i + 1e-3;
    c = {0.0, 0.0, -1};
    for (int j = threadIdx.x; j < n; j += SEL_THREADS)
      if (alive[j]) {
        const double tt1 = fmax(t1i, a1[j]), tt2 = fmin(t2i, a2[j]);
        const double wh = fmax(0.0, tt2 - tt1 + 1e-3);
        const double o = wh / (area_i + (a2[j] - a1[j] + 1e-3) - wh);
        if (o >= nms_overlap) {
          const RerankCand d = {ss[j], ps[j], j};
          if (better_pick(d, c)) c = d;
        }
        if (!(o <= nms_overlap)) alive[j] = 0;
      }
    const RerankCand pick = block_best<true>(c, smem);
    if (threadIdx.x == 0) op[npick] = pick.idx >= 0 ? pick.idx : i;
    ++npick;
  }
  if (threadIdx.x == 0) out_count[v] = npick;
}

}  // namespace echr

using namespace echr;

extern "C" int echr_select_topn(const float* scores, const float* mask, const int32_t* clip_offsets, int n_videos,
                                int K, int topn, float val_score_thres, int32_t* out_ind, int32_t* out_start,
                                int32_t* out_end, float* out_conf, int32_t* out_count, int cap, void* stream) {
  ECHR_REQUIRE(scores && clip_offsets && out_ind && out_start && out_end && out_conf && out_count,
               "echr_select_topn: null pointer");
  ECHR_REQUIRE(n_videos >= 0 && K >= 1 && topn >= 1 && cap >= 0, "echr_select_topn: bad shape");
  if (n_videos == 0) return ECHR_OK;
  select_topn_kernel<<<n_videos, SEL_THREADS, 0, (cudaStream_t)stream>>>(
      scores, mask, clip_offsets, K, topn, val_score_thres, out_ind, out_start, out_end, out_conf, out_count, cap);
  count_launch();
  ECHR_CUDA(cudaGetLastError());
  return ECHR_OK;
}

extern "C" int64_t echr_select_nms_proposals(int T, int K) {
  return T > 0 && K > 0 ? (int64_t)(T <= K + 1 ? (long long)T * (T - 1) / 2
                                                : (long long)K * (K + 1) / 2 + (long long)(T - 1 - K) * K)
                        : 0;
}

extern "C" int echr_select_nms(const float* scores, const int32_t* clip_offsets, int n_videos, int K, double overlap,
                               int topn, void* workspace, const int64_t* workspace_offsets, int64_t total_proposals,
                               int32_t* out_ind, int32_t* out_start, int32_t* out_end, float* out_score,
                               int32_t* out_pick, int32_t* out_count, int cap, void* stream) {
  ECHR_REQUIRE(scores && clip_offsets && workspace && workspace_offsets && out_ind && out_start && out_end &&
                   out_score && out_pick && out_count,
               "echr_select_nms: null pointer");
  ECHR_REQUIRE(n_videos >= 0 && K >= 1 && K <= 65535 && topn >= 1 && cap >= 0 && total_proposals >= 0,
               "echr_select_nms: bad shape");
  if (n_videos == 0) return ECHR_OK;
  uint32_t* key = reinterpret_cast<uint32_t*>(workspace);
  uint32_t* nk = key + total_proposals;
  select_nms_kernel<<<n_videos, SEL_THREADS, 0, (cudaStream_t)stream>>>(
      scores, clip_offsets, K, overlap, topn, key, nk, reinterpret_cast<const long long*>(workspace_offsets), out_ind,
      out_start, out_end, out_score, out_pick, out_count, cap);
  count_launch();
  ECHR_CUDA(cudaGetLastError());
  return ECHR_OK;
}

extern "C" int echr_rerank_nms(const double* t1, const double* t2, const double* prop_scores,
                               const double* sent_scores, const int32_t* offsets, int n_videos, double nms_overlap,
                               int topn, void* workspace, int32_t* out_pick, int32_t* out_count, int cap,
                               void* stream) {
  ECHR_REQUIRE(t1 && t2 && prop_scores && sent_scores && offsets && workspace && out_pick && out_count,
               "echr_rerank_nms: null pointer");
  ECHR_REQUIRE(n_videos >= 0 && topn >= 1 && cap >= 0, "echr_rerank_nms: bad shape");
  if (n_videos == 0) return ECHR_OK;
  const size_t smem = sizeof(RerankCand) * SEL_THREADS;
  rerank_nms_kernel<<<n_videos, SEL_THREADS, smem, (cudaStream_t)stream>>>(
      t1, t2, prop_scores, sent_scores, offsets, nms_overlap, topn, reinterpret_cast<uint8_t*>(workspace), out_pick,
      out_count, cap);
  count_launch();
  ECHR_CUDA(cudaGetLastError());
  return ECHR_OK;
}

// ---------------------------------------------------------------------------------------------
// echr_rerank_nms_static: the same kernel with its one-byte-per-proposal "alive" flags in a file-scope
// device array, for callers that must not allocate (no counterpart in the reference; see echr_b200.h).
// Every GPU that loads the library holds these ECHR_RERANK_STATIC_MAX bytes until the process exits.
static __device__ __align__(256) uint8_t g_rerank_alive[ECHR_RERANK_STATIC_MAX];

// In host code `g_rerank_alive` is only a stand-in for the symbol: its device address comes from
// cudaGetSymbolAddress, asked once per device with that device current.
static int rerank_static_scratch(uint8_t** out) {
  static uint8_t* addr[MAX_DEVICES] = {};
  int dev = -1;
  ECHR_CUDA(cudaGetDevice(&dev));
  ECHR_REQUIRE(dev >= 0 && dev < MAX_DEVICES, "echr_rerank_nms_static: device %d out of range", dev);
  if (addr[dev] == nullptr) {
    void* p = nullptr;
    ECHR_CUDA(cudaGetSymbolAddress(&p, g_rerank_alive));
    ECHR_REQUIRE(p != nullptr, "echr_rerank_nms_static: no device address for the scratch array");
    addr[dev] = static_cast<uint8_t*>(p);
  }
  *out = addr[dev];
  return ECHR_OK;
}

extern "C" int echr_rerank_nms_static(const double* t1, const double* t2, const double* prop_scores,
                                      const double* sent_scores, const int32_t* offsets, int n_videos,
                                      int64_t total_proposals, double nms_overlap, int topn, int32_t* out_pick,
                                      int32_t* out_count, int cap, void* stream) {
  ECHR_REQUIRE(t1 && t2 && prop_scores && sent_scores && offsets && out_pick && out_count,
               "echr_rerank_nms_static: null pointer");
  ECHR_REQUIRE(n_videos >= 0 && topn >= 1 && cap >= 0 && total_proposals >= 0, "echr_rerank_nms_static: bad shape");
  ECHR_REQUIRE(total_proposals <= (int64_t)ECHR_RERANK_STATIC_MAX,
               "echr_rerank_nms_static: %lld proposals exceed the built-in scratch of %d (ECHR_RERANK_STATIC_MAX); "
               "use echr_rerank_nms with a caller workspace",
               (long long)total_proposals, ECHR_RERANK_STATIC_MAX);
  if (n_videos == 0) return ECHR_OK;
  uint8_t* alive = nullptr;
  if (int rc = rerank_static_scratch(&alive); rc != ECHR_OK) return rc;
  // nothing to clear: the kernel sets the flags of every proposal it owns before reading them
  const size_t smem = sizeof(RerankCand) * SEL_THREADS;
  rerank_nms_kernel<<<n_videos, SEL_THREADS, smem, (cudaStream_t)stream>>>(
      t1, t2, prop_scores, sent_scores, offsets, nms_overlap, topn, alive, out_pick, out_count, cap);
  count_launch();
  ECHR_CUDA(cudaGetLastError());
  return ECHR_OK;
}
