// Context prologue and event relation encoder kernels (everything that runs once per batch):
//   split_planes        fp32 -> bf16 hi/lo planes, or one fp16 plane (tensor-core operand staging)
//   event_feats         mean-pooled clip feature | proposal state per event (CaptionGenerator.py:106-127)
//   relation_attn       multi-head relation attention with the pair weights (MA:122-175)
//   fill_video          LDA video context into the activation matrix (CaptionGenerator.py:87-104)
#include "engine.cuh"
#include "util.cuh"

namespace echr {

__global__ void split_planes_kernel(const float* __restrict__ src, long long rows, long long cols,
                                    long long ld_src, bf16* __restrict__ hi, bf16* __restrict__ lo,
                                    long long ld_dst, long long cols_dst, bool f16) {
  const long long total = rows * cols_dst;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cols_dst, c = i - r * cols_dst;
    const float x = (c < cols) ? src[r * ld_src + c] : 0.f;
    bf16 h, l;
    split16(x, f16, h, l);
    hi[r * ld_dst + c] = h;
    if (lo != nullptr) lo[r * ld_dst + c] = l;
  }
}

// 4 columns per thread: one 16-byte load, one 8-byte store per plane (rows of 4k columns, 16-byte aligned)
__global__ void split_planes_vec4_kernel(const float* __restrict__ src, long long rows, int cols, long long ld_src,
                                         bf16* __restrict__ hi, bf16* __restrict__ lo, long long ld_dst,
                                         int cols_dst, bool f16) {
  const int q_per_row = cols_dst >> 2;
  const long long total = rows * q_per_row;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / q_per_row;
    const int c = static_cast<int>(i - r * q_per_row) << 2;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < cols) x = __ldg(reinterpret_cast<const float4*>(src + r * ld_src + c));
    __align__(8) bf16 h[4], l[4];
    split16(x.x, f16, h[0], l[0]); split16(x.y, f16, h[1], l[1]);
    split16(x.z, f16, h[2], l[2]); split16(x.w, f16, h[3], l[3]);
    *reinterpret_cast<uint2*>(hi + r * ld_dst + c) = *reinterpret_cast<const uint2*>(h);
    if (lo != nullptr) *reinterpret_cast<uint2*>(lo + r * ld_dst + c) = *reinterpret_cast<const uint2*>(l);
  }
}

// f16: write one fp16 plane (lo must be null) instead of bf16 hi (+ lo)
cudaError_t launch_split(const float* src, int64_t rows, int64_t cols, int64_t ld_src, bf16* hi,
                         bf16* lo, int64_t ld_dst, int64_t cols_dst, bool f16, cudaStream_t st) {
  if (rows == 0 || cols_dst == 0) return cudaSuccess;
  if (f16 && lo != nullptr) return cudaErrorInvalidValue;
  const bool vec = (cols % 4 == 0) && (ld_src % 4 == 0) && (cols_dst % 4 == 0) && (ld_dst % 4 == 0) &&
                   ((reinterpret_cast<uintptr_t>(src) & 15) == 0) && ((reinterpret_cast<uintptr_t>(hi) & 7) == 0) &&
                   (lo == nullptr || (reinterpret_cast<uintptr_t>(lo) & 7) == 0);
  if (vec) {
    const long long total = rows * (cols_dst / 4);
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    split_planes_vec4_kernel<<<(int)blocks, 256, 0, st>>>(src, rows, (int)cols, ld_src, hi, lo, ld_dst, (int)cols_dst, f16);
    count_launch();
    return cudaGetLastError();
  }
  const long long total = rows * cols_dst;
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  split_planes_kernel<<<(int)blocks, 256, 0, st>>>(src, rows, cols, ld_src, hi, lo, ld_dst, cols_dst, f16);
  count_launch();
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// one block per event: 4 clip groups x 128 threads (one float4 of the 500 columns each) accumulate the
// span in parallel, then 256 threads write the 1024-wide feature row (mean clip feature | proposal state)
__global__ void __launch_bounds__(512)
event_feats_kernel(const float* __restrict__ c3d, const float* __restrict__ tap,
                   const int32_t* __restrict__ ev_start, const int32_t* __restrict__ ev_len,
                   bf16* __restrict__ hi, bf16* __restrict__ lo) {
  __shared__ __align__(16) float part[4][512];
  const int e = blockIdx.x;
  const int s = ev_start[e], len = ev_len[e];
  const int grp = threadIdx.x >> 7, t = threadIdx.x & 127;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (t < ECHR_VIDEO_DIM / 4) {
    const float* base = c3d + (long long)s * ECHR_VIDEO_DIM + 4 * t;
#pragma unroll 4
    for (int l = grp; l < len; l += 4) {
      const float4 c = __ldg(reinterpret_cast<const float4*>(base + (long long)l * ECHR_VIDEO_DIM));
      acc.x += c.x; acc.y += c.y; acc.z += c.z; acc.w += c.w;
    }
  }
  *reinterpret_cast<float4*>(&part[grp][4 * t]) = acc;
  __syncthreads();
  if (threadIdx.x < 256) {
    const int c0 = 4 * threadIdx.x;
    float v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = c0 + i;
      if (c < ECHR_VIDEO_DIM) v[i] = ((part[0][c] + part[1][c]) + (part[2][c] + part[3][c])) / (float)len;
      else if (c < ECHR_VIDEO_DIM + HID) v[i] = __ldg(tap + (long long)e * HID + (c - ECHR_VIDEO_DIM));  // gathered by ind at set_inputs
      else v[i] = 0.f;
    }
    __align__(8) bf16 h[4], l2[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) split16(v[i], lo == nullptr, h[i], l2[i]);
    *reinterpret_cast<uint2*>(hi + (long long)e * FEAT_PAD + c0) = *reinterpret_cast<const uint2*>(h);
    if (lo != nullptr) *reinterpret_cast<uint2*>(lo + (long long)e * FEAT_PAD + c0) = *reinterpret_cast<const uint2*>(l2);
  }
}

// tap_feats[ind_select_list] (CaptionGenerator.py:121): gather the proposal-encoder state of each
// event's last clip from the caller's [clips,512] tensor into the compact [events,512] workspace.
__global__ void __launch_bounds__(128)
gather_tap_kernel(const float* __restrict__ tap_src, const int32_t* __restrict__ ev_ind, float* __restrict__ dst) {
  const int e = blockIdx.x;
  const float4* src = reinterpret_cast<const float4*>(tap_src + (long long)ev_ind[e] * HID);
  reinterpret_cast<float4*>(dst + (long long)e * HID)[threadIdx.x] = __ldg(src + threadIdx.x);
}

cudaError_t launch_gather_tap(const echr_batch* b, const float* tap_src, cudaStream_t st) {
  if (b->n_events == 0) return cudaSuccess;
  gather_tap_kernel<<<b->n_events, 128, 0, st>>>(tap_src, b->ev_ind, b->tap);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_event_feats(const echr_batch* b, cudaStream_t st) {
  if (b->n_events == 0) return cudaSuccess;
  event_feats_kernel<<<b->n_events, 512, 0, st>>>(b->c3d, b->tap, b->ev_start, b->ev_len,
                                                  b->feat_pl.plane(0),
                                                  b->feat_pl.planes > 1 ? b->feat_pl.plane(1) : nullptr);
  count_launch();
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Relation attention (MA:122-175), one CTA per (video, head, block of REL_QB query events):
//   s[i,j] = w[i,j,g] * (Q_i,g . K_j,g) / sqrt(32)         fST0: multiplicative fusion (MA:148-149)
//   p[i,:] = softmax_j(s[i,:])                               (MA:160)
//   out[i, 32g+d] = sum_j p[i,j] * Y[j, 32g+d] + b[32g+d]    with Y = x * W_out^T  (grouped 1x1 conv,
//                                                            MA:98,167-175, re-associated)
// Keys are walked in chunks of REL_JC events whose K_g / Y_g rows sit in shared memory (rows padded to 33
// words: conflict-free for both the per-lane row walk of the scores and the per-lane column walk of the mix);
// one warp per query event, warp-shuffle softmax, running (max, sum, mix) per query kept in shared memory
// between chunks — so the number of events per video is not bounded by shared memory (the reference's default
// selection feeds up to topN = 1000 proposals, eval_utils.py:21).  Videos of up to REL_JC events take a single
// chunk: plain two-pass softmax.  The pair weights come head-major [16][n_pairs] from pair_mlp.cu.
constexpr int REL_WARPS = 8;
constexpr int REL_QB = 128;
constexpr int REL_JC = 256;

__global__ void __launch_bounds__(REL_WARPS * 32)
relation_attn_kernel(const float* __restrict__ qky, const float* __restrict__ pair_w, long long n_pairs,
                     const int32_t* __restrict__ ev_pair_off, const int32_t* __restrict__ vid_first,
                     const int32_t* __restrict__ vid_n, const float* __restrict__ b_out,
                     float* __restrict__ event_ctx, int jc, int qb) {
  extern __shared__ float rel_smem[];
  const int v = blockIdx.x, g = blockIdx.y;
  const int first = vid_first[v], n = vid_n[v];
  const int q0 = blockIdx.z * qb;
  if (q0 >= n) return;
  const int q1 = min(n, q0 + qb);
  float* Ks = rel_smem;                      // [jc][33]
  float* Ys = Ks + (size_t)jc * 33;          // [jc][33]
  float* Ss = Ys + (size_t)jc * 33;          // [REL_WARPS][jc]
  float* Qs = Ss + (size_t)REL_WARPS * jc;   // [REL_WARPS][32]
  float* Ms = Qs + REL_WARPS * 32;           // [qb] running max
  float* Ls = Ms + qb;                       // [qb] running sum
  float* As = Ls + qb;                       // [qb][32] running mix
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float scale = 1.0f / sqrtf(32.0f);
  float* S = Ss + (size_t)warp * jc;
  float* Q = Qs + warp * 32;
  for (int j0 = 0; j0 < n; j0 += jc) {
    const int jn = min(jc, n - j0);
    const bool first_chunk = j0 == 0, last_chunk = j0 + jc >= n;
    __syncthreads();  // every warp is done with the previous chunk
    for (int idx = threadIdx.x; idx < jn * 32; idx += blockDim.x) {
      const int j = idx >> 5, d = idx & 31;
      const float* row = qky + (long long)(first + j0 + j) * 1536 + 32 * g + d;
      Ks[j * 33 + d] = __ldg(row + 512);
      Ys[j * 33 + d] = __ldg(row + 1024);
    }
    __syncthreads();
    for (int i = q0 + warp; i < q1; i += REL_WARPS) {
      const int li = i - q0;
      Q[lane] = __ldg(qky + (long long)(first + i) * 1536 + 32 * g + lane);
      __syncwarp();
      const float* wrow = pair_w + (long long)g * n_pairs + ev_pair_off[first + i] + j0;
      float m_loc = -INFINITY;
      for (int j = lane; j < jn; j += 32) {
        float dot = 0.f;
#pragma unroll
        for (int d = 0; d < 32; ++d) dot = fmaf(Q[d], Ks[j * 33 + d], dot);
        const float s = __ldg(wrow + j) * (scale * dot);
        S[j] = s;
        m_loc = fmaxf(m_loc, s);
      }
      m_loc = warp_max(m_loc);
      const float m_old = first_chunk ? -INFINITY : Ms[li];
      const float m = fmaxf(m_old, m_loc);
      float sum = 0.f;
      for (int j = lane; j < jn; j += 32) {
        const float e = expf(S[j] - m);
        S[j] = e;
        sum += e;
      }
      sum = warp_sum(sum);
      __syncwarp();
      float acc = 0.f;
      if (!first_chunk) {
        const float corr = expf(m_old - m);
        acc = As[li * 32 + lane] * corr;
        sum += Ls[li] * corr;
      }
      for (int j = 0; j < jn; ++j) acc = fmaf(S[j], Ys[j * 33 + lane], acc);
      if (last_chunk) {
        event_ctx[(long long)(first + i) * HID + 32 * g + lane] = acc / sum + __ldg(b_out + 32 * g + lane);
      } else {
        As[li * 32 + lane] = acc;
        if (lane == 0) { Ms[li] = m; Ls[li] = sum; }
      }
      __syncwarp();
    }
  }
}

cudaError_t launch_relation_attn(const echr_batch* b, cudaStream_t st) {
  if (b->n_events == 0) return cudaSuccess;
  const int jc = b->max_n < REL_JC ? b->max_n : REL_JC;
  const int qb = b->max_n < REL_QB ? b->max_n : REL_QB;
  const size_t smem = ((size_t)jc * 66 + (size_t)REL_WARPS * jc + REL_WARPS * 32 + (size_t)qb * 34) * sizeof(float);
  static int attr[MAX_DEVICES] = {};
  if (smem > 48 * 1024) {
    cudaError_t e = set_max_dynamic_smem(relation_attn_kernel, (int)smem, attr);
    if (e != cudaSuccess) return e;
  }
  dim3 grid(b->n_videos, ECHR_HEADS, (b->max_n + qb - 1) / qb);
  relation_attn_kernel<<<grid, REL_WARPS * 32, smem, st>>>(b->qky, b->pair_w, (long long)b->n_pairs, b->ev_pair_off,
                                                          b->vid_first, b->vid_n, b->model->b_out, b->event_ctx, jc, qb);
  count_launch();
  return cudaGetLastError();
}

}  // namespace echr
