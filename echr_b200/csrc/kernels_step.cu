// Per-decode-step kernels that are not GEMMs:
//   attention_kernel   late-fusion attention over an event's clip window (OldModel_NEW.py:376-401)
//   decode_init        zero state, BOS embedding, constant context columns (OldModel_NEW.py:72-78,150-156)
//   finalize_greedy    merge the logit tiles' partials -> argmax / log-prob / next input (OldModel_NEW.py:158,175-183)
//   finalize_tf        same merge for the teacher-forced path (OldModel_NEW.py:98-130)
//   finalize_beam      beam selection + state reorder (spec: oracle/echr_oracle.py beam_search)
#include <type_traits>

#include "engine.cuh"
#include "util.cuh"

namespace echr {

// The attention score is sum_k alpha_k * tanh(P_lk + q_k).  With tanh(x) = 1 - 2 / (e^{2x} + 1) and
// e^{2(P+q)} = e^{2P} * e^{2q}, the tables hold EP = e^{2P} (written once per batch by the ctx2att
// GEMM epilogue) and EQ = e^{2q} (h2att GEMM epilogue), and a score costs per element one FMUL, one
// FADD, one MUFU.RCP and one FFMA:  score' = sum_k (-2 alpha_k) / (EP_lk * EQ_k + 1).  The dropped
// constant sum_k alpha_k is the same for every clip of the row, so the softmax is unchanged.  Both
// exponents are clamped to +-80 when the tables are written (ACT_EXP2X), so each factor is finite; their
// product may still overflow to +inf, which rcp.approx maps to 0, i.e. tanh = 1: the right limit.  The clamp
// itself is a documented bound of this library: |P| and |q| <= 40; only a pair with both beyond
// it and opposite signs would be evaluated differently from the reference.
__device__ __forceinline__ float rcp_fast(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

constexpr int ATT_THREADS = 512;        // latency-bound launches and long windows
constexpr int ATT_THREADS_SMALL = 256;  // many short windows: more CTAs per SM
constexpr int ATT_MAX_L = 1024;

// One CTA per decode row.  Pass 1: one warp per clip, lanes across the 512 hidden units (coalesced
// 16-byte loads), score_l = alpha . tanh(P_l + q).  Softmax over the window (the reference's
// softmax-then-mask-then-renormalise over the padded length, OldModel_NEW.py:394-397, is the same
// function).  Pass 2: threads across the 500 feature columns, att = sum_l w_l * clip_l.
template <bool BF16, int THREADS>
__global__ void __launch_bounds__(THREADS)
attention_kernel(const float* __restrict__ q, const void* __restrict__ proj_v,
                 const void* __restrict__ clip_v, long long ld_clip, const float* __restrict__ alpha,
                 const int32_t* __restrict__ row_start, const int32_t* __restrict__ row_len,
                 float* __restrict__ out_f32, bf16* __restrict__ out_hi, bf16* __restrict__ out_lo,
                 long long ld_bf, unsigned long long* trace, int trace_slot) {
  trace_begin(trace, trace_slot);
  __shared__ float sc[ATT_MAX_L];
  __shared__ __align__(16) float red[THREADS / 64][512];
  const int r = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_trigger();
  pdl_wait();  // q comes from the h2att GEMM launched just before; nothing but weights is safe earlier
  const int s0 = row_start[r];
  const int L = row_len[r];
  // q and alpha: lane owns hidden units [128*m + 4*lane, +4), m = 0..3
  float qv[16], av[16];
#pragma unroll
  for (int m = 0; m < 4; ++m) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(q + (long long)r * HID + 128 * m + 4 * lane));
    const float4 a = __ldg(reinterpret_cast<const float4*>(alpha + 128 * m + 4 * lane));
    qv[4 * m] = t.x; qv[4 * m + 1] = t.y; qv[4 * m + 2] = t.z; qv[4 * m + 3] = t.w;
    av[4 * m] = -2.f * a.x; av[4 * m + 1] = -2.f * a.y; av[4 * m + 2] = -2.f * a.z; av[4 * m + 3] = -2.f * a.w;
  }
  // two clips per warp iteration: both rows' loads are in flight before the first tanh
  auto load_proj = [&](int l, float* pv) {
    if (BF16) {
      const bf16* prow = reinterpret_cast<const bf16*>(proj_v) + (long long)(s0 + l) * HID;
#pragma unroll
      for (int m = 0; m < 4; ++m) {
        const uint2 u = __ldg(reinterpret_cast<const uint2*>(prow + 128 * m + 4 * lane));
        const __nv_bfloat162 b0 = *reinterpret_cast<const __nv_bfloat162*>(&u.x);
        const __nv_bfloat162 b1 = *reinterpret_cast<const __nv_bfloat162*>(&u.y);
        pv[4 * m] = __low2float(b0); pv[4 * m + 1] = __high2float(b0);
        pv[4 * m + 2] = __low2float(b1); pv[4 * m + 3] = __high2float(b1);
      }
    } else {
      const float* prow = reinterpret_cast<const float*>(proj_v) + (long long)(s0 + l) * HID;
#pragma unroll
      for (int m = 0; m < 4; ++m) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(prow + 128 * m + 4 * lane));
        pv[4 * m] = t.x; pv[4 * m + 1] = t.y; pv[4 * m + 2] = t.z; pv[4 * m + 3] = t.w;
      }
    }
  };
  constexpr int NW = THREADS / 32;
  for (int l = warp; l < L; l += 2 * NW) {
    const int l2 = l + NW;
    const bool two = l2 < L;
    float pa[16], pb2[16];
    load_proj(l, pa);
    load_proj(two ? l2 : l, pb2);
    float sa = 0.f, sb = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      sa = fmaf(av[i], rcp_fast(fmaf(pa[i], qv[i], 1.0f)), sa);
      sb = fmaf(av[i], rcp_fast(fmaf(pb2[i], qv[i], 1.0f)), sb);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      sa += __shfl_xor_sync(0xffffffffu, sa, o);
      sb += __shfl_xor_sync(0xffffffffu, sb, o);
    }
    if (lane == 0) {
      sc[l] = sa;
      if (two) sc[l2] = sb;
    }
  }
  __syncthreads();
  // softmax statistics (every warp computes them redundantly: L is a few hundred at most)
  float mx = -INFINITY;
  for (int l = lane; l < L; l += 32) mx = fmaxf(mx, sc[l]);
  mx = warp_max(mx);
  float sum = 0.f;
  for (int l = lane; l < L; l += 32) sum += expf(sc[l] - mx);
  sum = warp_sum(sum);
  __syncthreads();
  const float inv = 1.0f / sum;
  for (int l = threadIdx.x; l < L; l += THREADS) sc[l] = expf(sc[l] - mx) * inv;
  __syncthreads();
  // pass 2
  if (BF16) {
    // 64 threads per clip row (8 bf16 each), THREADS/64 clip groups
    const int grp = threadIdx.x >> 6, t = threadIdx.x & 63;
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    const bf16* cbase = reinterpret_cast<const bf16*>(clip_v) + (long long)s0 * ld_clip + 8 * t;
#pragma unroll 4
    for (int l = grp; l < L; l += THREADS / 64) {
      const uint4 u = __ldg(reinterpret_cast<const uint4*>(cbase + (long long)l * ld_clip));
      const float w = sc[l];
      const __nv_bfloat162* b = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        acc[2 * i] = fmaf(w, __low2float(b[i]), acc[2 * i]);
        acc[2 * i + 1] = fmaf(w, __high2float(b[i]), acc[2 * i + 1]);
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) red[grp][8 * t + i] = acc[i];
  } else {
    // 128 threads per clip row (4 floats each, 125 active), THREADS/128 clip groups
    const int grp = threadIdx.x >> 7, t = threadIdx.x & 127;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (t < ECHR_VIDEO_DIM / 4) {
      const float* cbase = reinterpret_cast<const float*>(clip_v) + (long long)s0 * ld_clip + 4 * t;
#pragma unroll 4
      for (int l = grp; l < L; l += THREADS / 128) {
        const float4 c = __ldg(reinterpret_cast<const float4*>(cbase + (long long)l * ld_clip));
        const float w = sc[l];
        acc.x = fmaf(w, c.x, acc.x); acc.y = fmaf(w, c.y, acc.y);
        acc.z = fmaf(w, c.z, acc.z); acc.w = fmaf(w, c.w, acc.w);
      }
    }
    *reinterpret_cast<float4*>(&red[grp][4 * t]) = acc;
  }
  __syncthreads();
  for (int c = threadIdx.x; c < ECHR_VIDEO_DIM; c += THREADS) {
    float v = 0.f;
#pragma unroll
    for (int g = 0; g < (BF16 ? THREADS / 64 : THREADS / 128); ++g) v += red[g][c];
    if (out_f32 != nullptr) out_f32[(long long)r * ECHR_VIDEO_DIM + c] = v;
    if (out_hi != nullptr) {
      bf16 h, l2;
      split16(v, out_lo == nullptr, h, l2);
      out_hi[(long long)r * ld_bf + c] = h;
      if (out_lo != nullptr) out_lo[(long long)r * ld_bf + c] = l2;
    }
  }
  trace_end(trace, trace_slot);
}

// Staged variant for latency-bound batches (few CTAs per SM): the event's whole window of projected
// rows and clip features is brought into shared memory by the bulk-copy engine (cp.async.bulk, two
// mbarriers) right at kernel start — windows are contiguous in HBM, so each is a handful of 1-D bulk
// copies — and both passes then run out of shared memory.  One global round trip instead of
// ~L/8 dependent ones.  Requires the window to fit: L * (proj row + clip row) <= ATT_STAGE_BYTES.
constexpr int ATT_STAGE_BYTES = 200 * 1024;
constexpr int ATT_BULK_CLIPS = 16;  // clips per bulk copy

template <bool BF16, int THREADS>
__global__ void __launch_bounds__(THREADS)
attention_staged_kernel(const float* __restrict__ q, const void* __restrict__ proj_v,
                        const void* __restrict__ clip_v, long long ld_clip,
                        const float* __restrict__ alpha, const int32_t* __restrict__ row_start,
                        const int32_t* __restrict__ row_len, float* __restrict__ out_f32,
                        bf16* __restrict__ out_hi, bf16* __restrict__ out_lo, long long ld_bf,
                        int max_len, unsigned long long* trace, int trace_slot) {
  trace_begin(trace, trace_slot);
  typedef typename std::conditional<BF16, bf16, float>::type T;
  constexpr int CLIP_ROW = BF16 ? CLIP_PAD : ECHR_VIDEO_DIM;  // elements per staged clip row
  extern __shared__ __align__(128) uint8_t att_smem[];
  __shared__ float sc[ATT_MAX_L];
  __shared__ __align__(16) float red[THREADS / 64][512];
  __shared__ __align__(8) uint64_t bars[2];
  T* sproj = reinterpret_cast<T*>(att_smem);                       // [max_len][512]
  T* sclip = sproj + (size_t)max_len * HID;                        // [max_len][CLIP_ROW]
  const int r = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_trigger();
  pdl_wait();
  const int s0 = row_start[r];
  const int L = row_len[r];
  const uint32_t bar_p = smem_u32(&bars[0]), bar_c = smem_u32(&bars[1]);
  if (threadIdx.x == 0) {
    mbar_init(bar_p, 1);
    mbar_init(bar_c, 1);
    fence_mbar_init();
    const uint32_t prow_b = HID * sizeof(T), crow_b = CLIP_ROW * sizeof(T);
    mbar_arrive_expect_tx(bar_p, (uint32_t)L * prow_b);
    mbar_arrive_expect_tx(bar_c, (uint32_t)L * crow_b);
    const uint8_t* psrc = reinterpret_cast<const uint8_t*>(proj_v) + (size_t)s0 * prow_b;
    const uint8_t* csrc = reinterpret_cast<const uint8_t*>(clip_v) + (size_t)s0 * (size_t)ld_clip * sizeof(T);
    for (int l = 0; l < L; l += ATT_BULK_CLIPS) {
      const uint32_t n = (uint32_t)min(ATT_BULK_CLIPS, L - l);
      bulk_load_1d(smem_u32(sproj) + l * prow_b, psrc + (size_t)l * prow_b, n * prow_b, bar_p);
    }
    for (int l = 0; l < L; l += ATT_BULK_CLIPS) {
      const uint32_t n = (uint32_t)min(ATT_BULK_CLIPS, L - l);
      bulk_load_1d(smem_u32(sclip) + l * crow_b, csrc + (size_t)l * crow_b, n * crow_b, bar_c);
    }
  }
  float qv[16], av[16];
#pragma unroll
  for (int m = 0; m < 4; ++m) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(q + (long long)r * HID + 128 * m + 4 * lane));
    const float4 a = __ldg(reinterpret_cast<const float4*>(alpha + 128 * m + 4 * lane));
    qv[4 * m] = t.x; qv[4 * m + 1] = t.y; qv[4 * m + 2] = t.z; qv[4 * m + 3] = t.w;
    av[4 * m] = -2.f * a.x; av[4 * m + 1] = -2.f * a.y; av[4 * m + 2] = -2.f * a.z; av[4 * m + 3] = -2.f * a.w;
  }
  __syncthreads();  // barrier init visible to every thread
  mbar_wait(bar_p, 0);
  for (int l = warp; l < L; l += THREADS / 32) {
    float s = 0.f;
    const T* prow = sproj + (size_t)l * HID;
#pragma unroll
    for (int m = 0; m < 4; ++m) {
      float p0, p1, p2, p3;
      if (BF16) {
        const uint2 u = *reinterpret_cast<const uint2*>(reinterpret_cast<const bf16*>(prow) + 128 * m + 4 * lane);
        const __nv_bfloat162 b0 = *reinterpret_cast<const __nv_bfloat162*>(&u.x);
        const __nv_bfloat162 b1 = *reinterpret_cast<const __nv_bfloat162*>(&u.y);
        p0 = __low2float(b0); p1 = __high2float(b0); p2 = __low2float(b1); p3 = __high2float(b1);
      } else {
        const float4 t = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(prow) + 128 * m + 4 * lane);
        p0 = t.x; p1 = t.y; p2 = t.z; p3 = t.w;
      }
      s = fmaf(av[4 * m], rcp_fast(fmaf(p0, qv[4 * m], 1.0f)), s);
      s = fmaf(av[4 * m + 1], rcp_fast(fmaf(p1, qv[4 * m + 1], 1.0f)), s);
      s = fmaf(av[4 * m + 2], rcp_fast(fmaf(p2, qv[4 * m + 2], 1.0f)), s);
      s = fmaf(av[4 * m + 3], rcp_fast(fmaf(p3, qv[4 * m + 3], 1.0f)), s);
    }
    s = warp_sum(s);
    if (lane == 0) sc[l] = s;
  }
  __syncthreads();
  float mx = -INFINITY;
  for (int l = lane; l < L; l += 32) mx = fmaxf(mx, sc[l]);
  mx = warp_max(mx);
  float sum = 0.f;
  for (int l = lane; l < L; l += 32) sum += expf(sc[l] - mx);
  sum = warp_sum(sum);
  __syncthreads();
  const float inv = 1.0f / sum;
  for (int l = threadIdx.x; l < L; l += THREADS) sc[l] = expf(sc[l] - mx) * inv;
  __syncthreads();
  mbar_wait(bar_c, 0);
  if (BF16) {
    const int grp = threadIdx.x >> 6, t = threadIdx.x & 63;
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    const bf16* cbase = reinterpret_cast<const bf16*>(sclip) + 8 * t;
    for (int l = grp; l < L; l += THREADS / 64) {
      const uint4 u = *reinterpret_cast<const uint4*>(cbase + (size_t)l * CLIP_ROW);
      const float w = sc[l];
      const __nv_bfloat162* b = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        acc[2 * i] = fmaf(w, __low2float(b[i]), acc[2 * i]);
        acc[2 * i + 1] = fmaf(w, __high2float(b[i]), acc[2 * i + 1]);
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) red[grp][8 * t + i] = acc[i];
  } else {
    const int grp = threadIdx.x >> 7, t = threadIdx.x & 127;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (t < ECHR_VIDEO_DIM / 4) {
      const float* cbase = reinterpret_cast<const float*>(sclip) + 4 * t;
      for (int l = grp; l < L; l += THREADS / 128) {
        const float4 c = *reinterpret_cast<const float4*>(cbase + (size_t)l * CLIP_ROW);
        const float w = sc[l];
        acc.x = fmaf(w, c.x, acc.x); acc.y = fmaf(w, c.y, acc.y);
        acc.z = fmaf(w, c.z, acc.z); acc.w = fmaf(w, c.w, acc.w);
      }
    }
    *reinterpret_cast<float4*>(&red[grp][4 * t]) = acc;
  }
  __syncthreads();
  for (int c = threadIdx.x; c < ECHR_VIDEO_DIM; c += THREADS) {
    float v = 0.f;
#pragma unroll
    for (int g = 0; g < (BF16 ? THREADS / 64 : THREADS / 128); ++g) v += red[g][c];
    if (out_f32 != nullptr) out_f32[(long long)r * ECHR_VIDEO_DIM + c] = v;
    if (out_hi != nullptr) {
      bf16 h, l2;
      split16(v, out_lo == nullptr, h, l2);
      out_hi[(long long)r * ld_bf + c] = h;
      if (out_lo != nullptr) out_lo[(long long)r * ld_bf + c] = l2;
    }
  }
  trace_end(trace, trace_slot);
}

// ---------------------------------------------------------------------------------------------
// Latency-regime variant (short windows, a few hundred rows): one CTA per decode row, every warp owns
// clips (warp, warp+NW, ...) in BOTH passes, so the clip features of the warp's first two clips are
// requested before the softmax and are in flight while the block agrees on the statistics.  A row
// costs two dependent global round trips (window descriptor -> everything else) and two block
// barriers; the streaming kernel above takes one round trip per 4 clips of pass 2 and four barriers.
template <bool BF16, int THREADS>
__global__ void __launch_bounds__(THREADS, 1024 / THREADS)
attention_v2_kernel(const float* __restrict__ q, const void* __restrict__ proj_v,
                    const void* __restrict__ clip_v, long long ld_clip, const float* __restrict__ alpha,
                    const int32_t* __restrict__ row_start, const int32_t* __restrict__ row_len,
                    float* __restrict__ out_f32, bf16* __restrict__ out_hi, bf16* __restrict__ out_lo,
                    long long ld_bf, unsigned long long* trace, int trace_slot) {
  trace_begin(trace, trace_slot);
  constexpr int NW = THREADS / 32;
  // shared memory is kept at 22 KB so that a CTA fits next to a resident GEMM CTA (~195 KB) of the
  // concurrent logit branch of the decode graph
  __shared__ float sc[ATT_MAX_L];
  __shared__ __align__(16) float red[NW / 2][512];
  __shared__ __align__(16) float s_alpha[HID];
  const int r = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < HID; i += THREADS) s_alpha[i] = -2.f * __ldg(alpha + i);
  pdl_trigger();
  pdl_wait();
  const int s0 = row_start[r];
  const int L = row_len[r];
  __syncthreads();  // s_alpha is read across warps in pass 1
  // lane owns hidden units [128*m + 4*lane, +4), m = 0..3, and clip columns float4 #(lane + 32*m)
  float qv[16];
#pragma unroll
  for (int m = 0; m < 4; ++m) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(q + (long long)r * HID + 128 * m + 4 * lane));
    qv[4 * m] = t.x; qv[4 * m + 1] = t.y; qv[4 * m + 2] = t.z; qv[4 * m + 3] = t.w;
  }
  auto load_row16 = [&](const void* base, long long ld, int cols, int l, float* v) {
    if (BF16) {
      const bf16* row = reinterpret_cast<const bf16*>(base) + (long long)(s0 + l) * ld;
#pragma unroll
      for (int m = 0; m < 4; ++m) {
        const uint2 u = __ldg(reinterpret_cast<const uint2*>(row + 128 * m + 4 * lane));
        const __nv_bfloat162 b0 = *reinterpret_cast<const __nv_bfloat162*>(&u.x);
        const __nv_bfloat162 b1 = *reinterpret_cast<const __nv_bfloat162*>(&u.y);
        v[4 * m] = __low2float(b0); v[4 * m + 1] = __high2float(b0);
        v[4 * m + 2] = __low2float(b1); v[4 * m + 3] = __high2float(b1);
      }
    } else {
      const float* row = reinterpret_cast<const float*>(base) + (long long)(s0 + l) * ld;
#pragma unroll
      for (int m = 0; m < 4; ++m) {
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
        if (128 * m + 4 * lane < cols) t = __ldg(reinterpret_cast<const float4*>(row + 128 * m + 4 * lane));
        v[4 * m] = t.x; v[4 * m + 1] = t.y; v[4 * m + 2] = t.z; v[4 * m + 3] = t.w;
      }
    }
  };
  const int clip_cols = BF16 ? CLIP_PAD : ECHR_VIDEO_DIM;
  // ---- pass 1: scores of this warp's clips, two at a time
  for (int l = warp; l < L; l += 2 * NW) {
    const int l2 = l + NW;
    const bool two = l2 < L;
    float pa[16], pb2[16];
    load_row16(proj_v, HID, HID, l, pa);
    load_row16(proj_v, HID, HID, two ? l2 : l, pb2);
    float sa = 0.f, sb = 0.f;
#pragma unroll
    for (int m = 0; m < 4; ++m) {
      const float4 a4 = *reinterpret_cast<const float4*>(s_alpha + 128 * m + 4 * lane);
      const float av[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        sa = fmaf(av[i], rcp_fast(fmaf(pa[4 * m + i], qv[4 * m + i], 1.0f)), sa);
        sb = fmaf(av[i], rcp_fast(fmaf(pb2[4 * m + i], qv[4 * m + i], 1.0f)), sb);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      sa += __shfl_xor_sync(0xffffffffu, sa, o);
      sb += __shfl_xor_sync(0xffffffffu, sb, o);
    }
    if (lane == 0) {
      sc[l] = sa;
      if (two) sc[l2] = sb;
    }
  }
  // ---- request the first two clip rows of this warp before the block agrees on the softmax
  float ca[16], cb[16];
  const bool has_a = warp < L, has_b = warp + NW < L;
  if (has_a) load_row16(clip_v, ld_clip, clip_cols, warp, ca);
  if (has_b) load_row16(clip_v, ld_clip, clip_cols, warp + NW, cb);
  __syncthreads();
  float mx = -INFINITY;
  for (int l = lane; l < L; l += 32) mx = fmaxf(mx, sc[l]);
  mx = warp_max(mx);
  float sum = 0.f;
  for (int l = lane; l < L; l += 32) sum += __expf(sc[l] - mx);
  sum = warp_sum(sum);
  const float inv = 1.0f / sum;
  // ---- pass 2: weighted sum of this warp's clips
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = 0.f;
  if (has_a) {
    const float w = __expf(sc[warp] - mx) * inv;
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = w * ca[i];
  }
  if (has_b) {
    const float w = __expf(sc[warp + NW] - mx) * inv;
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = fmaf(w, cb[i], acc[i]);
  }
  for (int l = warp + 2 * NW; l < L; l += 2 * NW) {
    const int l2 = l + NW;
    const bool two = l2 < L;
    load_row16(clip_v, ld_clip, clip_cols, l, ca);
    load_row16(clip_v, ld_clip, clip_cols, two ? l2 : l, cb);
    const float wa = __expf(sc[l] - mx) * inv;
    const float wb = two ? __expf(sc[l2] - mx) * inv : 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = fmaf(wa, ca[i], fmaf(wb, cb[i], acc[i]));
  }
  // cross-warp sum in two rounds (upper half of the warps hands its partial sums to the lower half)
  if (warp >= NW / 2) {
#pragma unroll
    for (int m = 0; m < 4; ++m)
      *reinterpret_cast<float4*>(&red[warp - NW / 2][128 * m + 4 * lane]) =
          make_float4(acc[4 * m], acc[4 * m + 1], acc[4 * m + 2], acc[4 * m + 3]);
  }
  __syncthreads();
  if (warp < NW / 2) {
#pragma unroll
    for (int m = 0; m < 4; ++m) {
      float4* p = reinterpret_cast<float4*>(&red[warp][128 * m + 4 * lane]);
      float4 t = make_float4(acc[4 * m], acc[4 * m + 1], acc[4 * m + 2], acc[4 * m + 3]);
      if (warp + NW / 2 < L) {  // the partner warp owned at least one clip
        const float4 o = *p;
        t.x += o.x; t.y += o.y; t.z += o.z; t.w += o.w;
      }
      *p = t;
    }
  }
  __syncthreads();
  const int nw_used = L < NW / 2 ? L : NW / 2;  // warps beyond the window hold zeros
  for (int c = threadIdx.x; c < ECHR_VIDEO_DIM; c += THREADS) {
    float v = 0.f;
    for (int g = 0; g < nw_used; ++g) v += red[g][c];
    if (out_f32 != nullptr) out_f32[(long long)r * ECHR_VIDEO_DIM + c] = v;
    if (out_hi != nullptr) {
      bf16 h, l2;
      split16(v, out_lo == nullptr, h, l2);
      out_hi[(long long)r * ld_bf + c] = h;
      if (out_lo != nullptr) out_lo[(long long)r * ld_bf + c] = l2;
    }
  }
  trace_end(trace, trace_slot);
}

// max_len: largest window of the launch (0 = unknown).  Variant choice (measured, profiles/r01_notes.md):
//   * staged kernel (one CTA per SM, ~200 KB of shared memory): every window fits and the launch is a single
//     wave (rows <= 148);
//   * attention_v2_kernel (warps own clips, clip rows requested before the softmax barrier, 256 threads): every
//     other launch whose longest window is known — 8.1 vs 10.1 us at 251 rows, 55 vs 71 us at 4072 rows, 82 vs
//     101 us at 3200 rows x 256-clip windows against the streaming kernel;
//   * streaming attention_kernel: max_len unknown (stand-alone echr_attention_step calls), or forced.
// ECHR_ATT_STAGED=1 / ECHR_ATT_STREAM=1 / ECHR_ATT_V2_ROWS=0 force one or the other.
// rows up to which the latency-regime kernel is used (ECHR_ATT_V2_ROWS; 0 disables it)
static int att_v2_rows() {
  static int v = -1;
  if (v < 0) {
    const char* e = knob("ECHR_ATT_V2_ROWS");
    v = e ? atoi(e) : (1 << 30);
  }
  return v;
}
// longest window the latency kernel takes (ECHR_ATT_V2_MAXLEN)
static int att_v2_maxlen() {
  static int v = -1;
  if (v < 0) {
    const char* e = knob("ECHR_ATT_V2_MAXLEN");
    v = e ? atoi(e) : ATT_MAX_L;
  }
  return v;
}

cudaError_t launch_attention(const float* q, const void* proj, const void* clip, int64_t ld_clip,
                             int clip_bf16, const float* alpha, const int32_t* row_start,
                             const int32_t* row_len, int64_t rows, float* out_f32, bf16* out_hi,
                             bf16* out_lo, int64_t ld_bf, cudaStream_t st, int max_len, unsigned long long* trace,
                             int trace_slot) {
  if (rows == 0) return cudaSuccess;
  cudaError_t e;
  const size_t el = clip_bf16 ? 2 : 4;
  const size_t clip_row = clip_bf16 ? CLIP_PAD : ECHR_VIDEO_DIM;
  const size_t stage = (size_t)max_len * (HID + clip_row) * el;
  // staged rows must be contiguous: fp32 clips have ld 500, bf16 clips ld 512
  const bool fits = max_len > 0 && max_len <= ATT_MAX_L && stage <= ATT_STAGE_BYTES && (size_t)ld_clip == clip_row;
  const bool staged = fits && knob("ECHR_ATT_STREAM") == nullptr && (rows <= 148 || knob("ECHR_ATT_STAGED") != nullptr);
  if (staged) {
    static size_t attr[2] = {0, 0};
    if (stage > attr[clip_bf16 ? 1 : 0]) {
      e = clip_bf16 ? cudaFuncSetAttribute(attention_staged_kernel<true, ATT_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_STAGE_BYTES)
                    : cudaFuncSetAttribute(attention_staged_kernel<false, ATT_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_STAGE_BYTES);
      if (e != cudaSuccess) return e;
      attr[clip_bf16 ? 1 : 0] = ATT_STAGE_BYTES;
    }
    if (clip_bf16)
      e = launch_pdl(attention_staged_kernel<true, ATT_THREADS>, dim3((unsigned)rows), dim3(ATT_THREADS), stage, st, q, proj, clip,
                     (long long)ld_clip, alpha, row_start, row_len, out_f32, out_hi, out_lo, (long long)ld_bf, max_len, trace, trace_slot);
    else
      e = launch_pdl(attention_staged_kernel<false, ATT_THREADS>, dim3((unsigned)rows), dim3(ATT_THREADS), stage, st, q, proj, clip,
                     (long long)ld_clip, alpha, row_start, row_len, out_f32, out_hi, out_lo, (long long)ld_bf, max_len, trace, trace_slot);
  } else if (att_v2_rows() > 0 && rows <= att_v2_rows() && max_len > 0 && max_len <= att_v2_maxlen()) {
    // latency regime: short windows.  256-thread CTAs (16 K registers, 13 KB) also fit next to a resident
    // CTA of the concurrent logit GEMM, so the launch is one wave even while that GEMM holds 90 SMs.
    static const int v2_threads = knob("ECHR_ATT_V2_THREADS") ? atoi(knob("ECHR_ATT_V2_THREADS")) : 256;
    if (v2_threads == 256) {
      if (clip_bf16)
        e = launch_pdl(attention_v2_kernel<true, 256>, dim3((unsigned)rows), dim3(256), 0, st, q, proj, clip,
                       (long long)ld_clip, alpha, row_start, row_len, out_f32, out_hi, out_lo, (long long)ld_bf, trace, trace_slot);
      else
        e = launch_pdl(attention_v2_kernel<false, 256>, dim3((unsigned)rows), dim3(256), 0, st, q, proj, clip,
                       (long long)ld_clip, alpha, row_start, row_len, out_f32, out_hi, out_lo, (long long)ld_bf, trace, trace_slot);
    } else if (clip_bf16)
      e = launch_pdl(attention_v2_kernel<true, ATT_THREADS>, dim3((unsigned)rows), dim3(ATT_THREADS), 0, st, q, proj, clip,
                     (long long)ld_clip, alpha, row_start, row_len, out_f32, out_hi, out_lo, (long long)ld_bf, trace, trace_slot);
    else
      e = launch_pdl(attention_v2_kernel<false, ATT_THREADS>, dim3((unsigned)rows), dim3(ATT_THREADS), 0, st, q, proj, clip,
                     (long long)ld_clip, alpha, row_start, row_len, out_f32, out_hi, out_lo, (long long)ld_bf, trace, trace_slot);
  } else {
    // many short windows (throughput regime): 256-thread CTAs pack more rows per SM; otherwise 512
    const bool small = rows > 148 * 4 && max_len > 0 && max_len <= 64;
    if (clip_bf16 && small)
      e = launch_pdl(attention_kernel<true, ATT_THREADS_SMALL>, dim3((unsigned)rows), dim3(ATT_THREADS_SMALL), 0, st, q, proj,
                     clip, (long long)ld_clip, alpha, row_start, row_len, out_f32, out_hi, out_lo, (long long)ld_bf, trace, trace_slot);
    else if (clip_bf16)
      e = launch_pdl(attention_kernel<true, ATT_THREADS>, dim3((unsigned)rows), dim3(ATT_THREADS), 0, st, q, proj, clip,
                     (long long)ld_clip, alpha, row_start, row_len, out_f32, out_hi, out_lo, (long long)ld_bf, trace, trace_slot);
    else if (small)
      e = launch_pdl(attention_kernel<false, ATT_THREADS_SMALL>, dim3((unsigned)rows), dim3(ATT_THREADS_SMALL), 0, st, q, proj,
                     clip, (long long)ld_clip, alpha, row_start, row_len, out_f32, out_hi, out_lo, (long long)ld_bf, trace, trace_slot);
    else
      e = launch_pdl(attention_kernel<false, ATT_THREADS>, dim3((unsigned)rows), dim3(ATT_THREADS), 0, st, q, proj, clip,
                     (long long)ld_clip, alpha, row_start, row_len, out_f32, out_hi, out_lo, (long long)ld_bf, trace, trace_slot);
  }
  count_launch();
  return e;
}

// ---------------------------------------------------------------------------------------------
// 16-bit, bulk-copy staged attention: design notes in attention_f16.cuh.
template <int ATR_WARPS>
__global__ void __launch_bounds__(ATR_WARPS * 32, 24 / ATR_WARPS)
attention_f16_kernel(const __grid_constant__ AttF16Params P, unsigned long long* trace, int trace_slot) {
  extern __shared__ __align__(128) uint8_t atr_smem[];
  trace_begin(trace, trace_slot);
  constexpr float L2E = 1.4426950408889634f;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x;
  const uint32_t ring = smem_u32(atr_smem) + warp * ATR_WARP_BYTES;   // this warp's slots
  uint8_t* tail = atr_smem + ATR_WARPS * ATR_WARP_BYTES;
  const uint32_t bars = smem_u32(tail) + warp * (ATR_SLOTS * 8);                     // this warp's mbarriers
  float* s_m = reinterpret_cast<float*>(tail + 384);   // [8] per-warp running max
  float* s_l = reinterpret_cast<float*>(tail + 416);   // [8] per-warp running sum
  if (lane == 0) {
    for (int i = 0; i < ATR_SLOTS; ++i) mbar_init(bars + 8u * i, 1);
    fence_mbar_init();
  }
  __syncwarp();
  pdl_trigger();
  pdl_wait();  // e^{2q} comes from the h2att GEMM launched just before
  const int s0 = __ldg(P.row_start + r);
  const int L = __ldg(P.row_len + r);
  const int n_mine = L > warp ? (L - warp + ATR_WARPS - 1) / ATR_WARPS : 0;  // clips warp, warp + 8, ...
  // this warp's clips: table / clip rows (s0 + warp + i W), i = 0 .. n_mine - 1
  const __nv_bfloat16* tbase = P.tab + static_cast<long long>(s0 + warp) * 1024;
  const float* sbase = P.proj_scale + s0 + warp;
  auto issue = [&](int i, int slot) {  // lane 0: bring this warp's i-th clip into `slot`
    const uint32_t dst = ring + slot * ATR_SLOT_BYTES;
    const uint32_t bar = bars + 8u * slot;
    mbar_arrive_expect_tx(bar, ATR_SLOT_BYTES);
    bulk_load_1d(dst, tbase + static_cast<long long>(i) * (ATR_WARPS * 1024), ATR_SLOT_BYTES, bar);
  };
  if (lane == 0)
    for (int i = 0; i < ATR_SLOTS && i < n_mine; ++i) issue(i, i);
  // (everything below is loaded while the first copies are in flight: the row's dependent chain is
  //  window descriptor -> copies -> compute -> merge, nothing else)
  // lane's hidden units / clip columns: [8 lane + 256 m, +8), m = 0, 1 (16-byte chunks lane and lane + 32 of a row)
  float av[16];
#pragma unroll
  for (int m = 0; m < 2; ++m) {
    const float4 a0 = __ldg(reinterpret_cast<const float4*>(P.alpha + 256 * m + 8 * lane));
    const float4 a1 = __ldg(reinterpret_cast<const float4*>(P.alpha + 256 * m + 8 * lane + 4));
    av[8 * m + 0] = -2.f * a0.x; av[8 * m + 1] = -2.f * a0.y; av[8 * m + 2] = -2.f * a0.z; av[8 * m + 3] = -2.f * a0.w;
    av[8 * m + 4] = -2.f * a1.x; av[8 * m + 5] = -2.f * a1.y; av[8 * m + 6] = -2.f * a1.z; av[8 * m + 7] = -2.f * a1.w;
  }
  float qe[16];
#pragma unroll
  for (int m = 0; m < 2; ++m) {
    const float* qp = P.q_exp + static_cast<long long>(r) * 512 + 256 * m + 8 * lane;
    const float4 q0 = __ldg(reinterpret_cast<const float4*>(qp));
    const float4 q1 = __ldg(reinterpret_cast<const float4*>(qp + 4));
    qe[8 * m + 0] = q0.x; qe[8 * m + 1] = q0.y; qe[8 * m + 2] = q0.z; qe[8 * m + 3] = q0.w;
    qe[8 * m + 4] = q1.x; qe[8 * m + 5] = q1.y; qe[8 * m + 6] = q1.z; qe[8 * m + 7] = q1.w;
  }
  // 1 / S of the warp's next clip (S = e^{2 rowmax P}): e^{2x} + 1 = S (E q + 1/S), so the per-clip scale enters the
  // FFMA that forms the denominator and multiplies the clip's score once, instead of every element
  auto rcp_fast = [](float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; };
  float is_next = n_mine > 0 ? rcp_fast(__ldg(sbase)) : 0.f;
  float mx = -INFINITY, sum = 0.f;
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = 0.f;
  // slots are walked round robin, so the loop is unrolled by the ring depth: slot and barrier addresses are constants
  for (int base = 0; base < n_mine; base += ATR_SLOTS) {
    const uint32_t parity = (base / ATR_SLOTS) & 1u;
#pragma unroll
    for (int slot = 0; slot < ATR_SLOTS; ++slot) {
      const int i = base + slot;
      if (i >= n_mine) break;
      const float inv_s = is_next;
      if (i + 1 < n_mine) is_next = rcp_fast(__ldg(sbase + (i + 1) * ATR_WARPS));
      mbar_wait_hot(bars + 8u * slot, parity);
      const uint32_t sp = ring + slot * ATR_SLOT_BYTES;
      uint4 pu[2], cu[2];
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(pu[m].x), "=r"(pu[m].y), "=r"(pu[m].z), "=r"(pu[m].w)
                     : "r"(sp + 16u * (lane + 32 * m)));
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(cu[m].x), "=r"(cu[m].y), "=r"(cu[m].z), "=r"(cu[m].w)
                     : "r"(sp + 1024u + 16u * (lane + 32 * m)));
      }
      float s = 0.f;
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        const uint32_t w4[4] = {pu[m].x, pu[m].y, pu[m].z, pu[m].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 e2 = f16x2_value(w4[j]);
          s = fmaf(av[8 * m + 2 * j], rcp_fast(fmaf(e2.x, qe[8 * m + 2 * j], inv_s)), s);
          s = fmaf(av[8 * m + 2 * j + 1], rcp_fast(fmaf(e2.y, qe[8 * m + 2 * j + 1], inv_s)), s);
        }
      }
      s = warp_sum(s) * inv_s;
      if (s > mx) {  // (warp-uniform) the running maximum moves: rescale what has been accumulated — O(log n) times per row
        float corr;
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(corr) : "f"((mx - s) * L2E));  // ex2(-inf) = 0 on the warp's first clip
        sum *= corr;
#pragma unroll
        for (int i2 = 0; i2 < 16; ++i2) acc[i2] *= corr;
        mx = s;
      }
      float p;
      asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p) : "f"((s - mx) * L2E));
      sum += p;
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        const uint32_t w4[4] = {cu[m].x, cu[m].y, cu[m].z, cu[m].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 c2 = f16x2_value(w4[j]);
          acc[8 * m + 2 * j] = fmaf(p, c2.x, acc[8 * m + 2 * j]);
          acc[8 * m + 2 * j + 1] = fmaf(p, c2.y, acc[8 * m + 2 * j + 1]);
        }
      }
      // refill the slot: every lane's reads of it have been consumed by the arithmetic above (and __syncwarp orders
      // the lanes), so the copy can only land after them
      __syncwarp();
      if (lane == 0 && i + ATR_SLOTS < n_mine) issue(i + ATR_SLOTS, slot);
    }
  }
  // ---- merge the eight warps' (max, sum, weighted sum): att = sum_w e^{m_w - M} acc_w / sum_w e^{m_w - M} sum_w
  if (lane == 0) { s_m[warp] = mx; s_l[warp] = sum; }
  __syncthreads();  // also: every warp is past its last slot and no copy is in flight -> the rings are free
  float M = -INFINITY;
#pragma unroll
  for (int w = 0; w < ATR_WARPS; ++w) M = fmaxf(M, s_m[w]);
  float total = 0.f;
#pragma unroll
  for (int w = 0; w < ATR_WARPS; ++w) total += s_m[w] == -INFINITY ? 0.f : s_l[w] * exp2f((s_m[w] - M) * L2E);
  const float my = mx == -INFINITY ? 0.f : exp2f((mx - M) * L2E);
  float* red = reinterpret_cast<float*>(atr_smem);  // [W][512] over the first quarter of the rings
#pragma unroll
  for (int m = 0; m < 2; ++m) {
    float* d = red + warp * 512 + 256 * m + 8 * lane;
    *reinterpret_cast<float4*>(d) = make_float4(acc[8 * m] * my, acc[8 * m + 1] * my, acc[8 * m + 2] * my, acc[8 * m + 3] * my);
    *reinterpret_cast<float4*>(d + 4) = make_float4(acc[8 * m + 4] * my, acc[8 * m + 5] * my, acc[8 * m + 6] * my, acc[8 * m + 7] * my);
  }
  __syncthreads();
  const float inv = total > 0.f ? 1.0f / total : 0.f;  // (an empty window gives a zero feature)
  for (int c0 = 2 * threadIdx.x; c0 < 512; c0 += 64 * ATR_WARPS) {  // two columns per thread and pass
    float v0 = 0.f, v1 = 0.f;
#pragma unroll
    for (int w = 0; w < ATR_WARPS; ++w) {
      const float2 t = *reinterpret_cast<const float2*>(red + w * 512 + c0);
      v0 += t.x;
      v1 += t.y;
    }
    v0 *= inv;
    v1 *= inv;
    if (P.out != nullptr) {
      __align__(4) __nv_bfloat16 h[2] = {f16_bits(v0), f16_bits(v1)};
      *reinterpret_cast<uint32_t*>(P.out + static_cast<long long>(r) * P.ld_out + c0) = *reinterpret_cast<const uint32_t*>(h);
    }
    if (P.out_f32 != nullptr && c0 < 500) {
      P.out_f32[static_cast<long long>(r) * 500 + c0] = v0;
      P.out_f32[static_cast<long long>(r) * 500 + c0 + 1] = v1;
    }
  }
  trace_end(trace, trace_slot);
}

template <int W>
static cudaError_t launch_attention_f16_w(const AttF16Params& P, cudaStream_t st, unsigned long long* trace, int trace_slot) {
  static int attr[MAX_DEVICES] = {};
  cudaError_t e = set_max_dynamic_smem(attention_f16_kernel<W>, atr_smem(W), attr);
  if (e != cudaSuccess) return e;
  return launch_pdl(attention_f16_kernel<W>, dim3((unsigned)P.rows), dim3(W * 32), (size_t)atr_smem(W), st, P, trace, trace_slot);
}

// mean_len: average window length of the launch (0 = unknown).  See attention_f16.cuh for the choice of W.
cudaError_t launch_attention_f16(const AttF16Params& P, int mean_len, cudaStream_t st, unsigned long long* trace, int trace_slot) {
  if (P.rows == 0) return cudaSuccess;
  cudaError_t e;
  const char* force = knob("ECHR_ATT_W");  // (debug builds only)
  if (force != nullptr) {
    const int w = atoi(force);
    e = w == 2 ? launch_attention_f16_w<2>(P, st, trace, trace_slot) : w == 4 ? launch_attention_f16_w<4>(P, st, trace, trace_slot)
                                                                              : launch_attention_f16_w<8>(P, st, trace, trace_slot);
    count_launch();
    return e;
  }
  if (P.rows <= 2 * num_sms() || mean_len <= 0) e = launch_attention_f16_w<8>(P, st, trace, trace_slot);
  else if (mean_len > 16) e = launch_attention_f16_w<4>(P, st, trace, trace_slot);
  else e = launch_attention_f16_w<2>(P, st, trace, trace_slot);
  count_launch();
  return e;
}

// Attention table of the single-plane precision (attention_f16.cuh): one warp per clip row,
// P fp32 [clips,512] -> E = e^{2 (P - rowmax)} as fp16 (row stride ld) and S = e^{2 rowmax} as fp32.
__global__ void __launch_bounds__(256)
attention_table_kernel(const float* __restrict__ proj, int clips, bf16* __restrict__ table, long long ld, float* __restrict__ scale) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int l = blockIdx.x * 8 + warp;
  if (l >= clips) return;
  const float* row = proj + static_cast<long long>(l) * HID;
  float v[16];
#pragma unroll
  for (int m = 0; m < 4; ++m) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(row + 128 * m + 4 * lane));
    v[4 * m] = t.x; v[4 * m + 1] = t.y; v[4 * m + 2] = t.z; v[4 * m + 3] = t.w;
  }
  float mx = v[0];
#pragma unroll
  for (int i = 1; i < 16; ++i) mx = fmaxf(mx, v[i]);
  mx = warp_max(mx);
#pragma unroll
  for (int m = 0; m < 4; ++m) {
    __align__(8) bf16 h[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = f16_bits(__expf(2.0f * (v[4 * m + i] - mx)));
    *reinterpret_cast<uint2*>(table + static_cast<long long>(l) * ld + 128 * m + 4 * lane) = *reinterpret_cast<const uint2*>(h);
  }
  if (lane == 0) scale[l] = __expf(2.0f * fminf(mx, 43.0f));  // (finite for any finite projection)
}

cudaError_t launch_attention_table(const float* proj, int clips, bf16* table, long long ld, float* scale, cudaStream_t st) {
  if (clips == 0) return cudaSuccess;
  attention_table_kernel<<<(clips + 7) / 8, 256, 0, st>>>(proj, clips, table, ld, scale);
  count_launch();
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// decode_init: one CTA per decode row (row = event * beam + slot).  The activation buffers and
// c-state were zeroed by memset nodes before this kernel.
__global__ void __launch_bounds__(128)
decode_init_kernel(int beam, int a_planes, const float* __restrict__ event_ctx,
                   const float* __restrict__ lda, const int32_t* __restrict__ ev_video,
                   const int32_t* __restrict__ ev_start, const int32_t* __restrict__ ev_len,
                   bf16* __restrict__ act0,
                   bf16* __restrict__ act1, long long act_plane, int32_t* __restrict__ row_start,
                   int32_t* __restrict__ row_len, int32_t* __restrict__ unfinished, int32_t* __restrict__ row_tok,
                   unsigned long long* __restrict__ key0, unsigned long long* __restrict__ key1,
                   float* __restrict__ beam_score, int32_t* __restrict__ beam_live,
                   float* __restrict__ best_score) {
  const int r = blockIdx.x;
  const int e = r / beam, slot = r - e * beam;
  const int v = ev_video[e];
  bf16* rows[2] = {act0 + (long long)r * ACT_LD, act1 + (long long)r * ACT_LD};
  for (int c = threadIdx.x; c < HID + VID_PAD; c += blockDim.x) {
    float x;
    int col;
    if (c < HID) {
      x = __ldg(event_ctx + (long long)e * HID + c);
      col = ACT_EV + c;
    } else {
      const int d = c - HID;
      x = d < ECHR_LDA_DIM ? __ldg(lda + (long long)v * ECHR_LDA_DIM + d) : 0.f;
      col = ACT_VID + d;
    }
    bf16 h, l;
    split16(x, a_planes == 1, h, l);
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      rows[b][col] = h;
      if (a_planes > 1) rows[b][act_plane + col] = l;
    }
  }
  if (threadIdx.x == 0) {
    row_tok[r] = 0;  // BOS (token 0)
    key0[(long long)r * KEY_STRIDE] = 0ull;                           // running arg-max of the logits of step 0
    key1[(long long)r * KEY_STRIDE] = pack_argmax_key(0.f, 0);        // "previous arg-max" read by step 0: BOS
    row_start[r] = ev_start[e];
    row_len[r] = ev_len[e];
    unfinished[r] = 1;
    beam_score[r] = 0.f;
    beam_live[r] = (slot == 0) ? 1 : 0;
    if (slot == 0) best_score[e] = -INFINITY;
  }
}

cudaError_t launch_decode_init(const echr_batch* b, int beam, cudaStream_t st) {
  const int rows = b->n_events * beam;
  if (rows == 0) return cudaSuccess;
  const echr_model* m = b->model;
  decode_init_kernel<<<rows, 128, 0, st>>>(
      beam, (int)b->act[0].planes, b->event_ctx, b->lda, b->ev_video, b->ev_start, b->ev_len,
      b->act[0].p, b->act[1].p, b->act[0].plane_stride(),
      b->row_start, b->row_len, b->unfinished, b->row_tok, b->argmax_key[0], b->argmax_key[1], b->beam_score,
      b->beam_live, b->best_score);
  count_launch();
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Merge of the logit GEMM's per-tile partial records of one row, by one warp.
// Record layout (LogitRec<TOPK>): [0] sum exp(x - v0), [1..TOPK] values desc, [1+TOPK..] indices.
struct RowStat {
  float max_val;   // largest logit of the row
  int max_idx;     // its column (lowest column among equals, like torch.max)
  float log_sum;   // log sum_j exp(x_j - max_val)   => log-prob of column j = x_j - max_val - log_sum
};

template <int TOPK>
__device__ __forceinline__ RowStat merge_row_stats(const float* __restrict__ rec_row, int n_tiles,
                                                   int lane) {
  constexpr int W = LogitRec<TOPK>::WORDS;
  float bv = -INFINITY;
  int bi = 0x7fffffff;
  for (int t = lane; t < n_tiles; t += 32) {
    const float v = rec_row[t * W + 1];
    const int i = __float_as_int(rec_row[t * W + 1 + TOPK]);
    if (v > bv || (v == bv && i < bi)) { bv = v; bi = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
  }
  float sum = 0.f;
  for (int t = lane; t < n_tiles; t += 32)
    sum += rec_row[t * W] * expf(rec_row[t * W + 1] - bv);
  sum = warp_sum(sum);
  RowStat s;
  s.max_val = bv;
  s.max_idx = bi;
  s.log_sum = logf(sum);
  return s;
}

constexpr int FIN_WARPS = 4;

// Greedy bookkeeping of OldModel.sample (OldModel_NEW.py:154-183): step `t` here is the
// reference's iteration t+1 — token = argmax of this step's log-probs, `unfinished` is and-ed with
// token > 0, the stored token is masked, the stored log-prob is NOT, and the raw argmax is what
// gets embedded for the next step.
__global__ void __launch_bounds__(FIN_WARPS * 32)
finalize_greedy_kernel(int rows, int n_tiles, int step, int seq_len, const float* __restrict__ partial,
                       int32_t* __restrict__ row_tok, int32_t* __restrict__ unfinished,
                       int32_t* __restrict__ tokens, float* __restrict__ logp, unsigned long long* __restrict__ key_next,
                       unsigned long long* trace, int trace_slot) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * FIN_WARPS + warp;
  trace_begin(trace, trace_slot);
  pdl_trigger();
  pdl_wait();
  if (r >= rows) return;
  const RowStat s = merge_row_stats<1>(partial + (long long)r * n_tiles * LogitRec<1>::WORDS, n_tiles, lane);
  const int tok = s.max_idx;
  if (lane == 0) {
    const int unf = (step == 0 ? 1 : unfinished[r]) & (tok > 0 ? 1 : 0);
    unfinished[r] = unf;
    tokens[(long long)r * seq_len + step] = unf ? tok : 0;
    logp[(long long)r * seq_len + step] = -s.log_sum;
    row_tok[r] = tok;  // the raw argmax is the next step's input
    // the arg-max keys the NEXT logit launch accumulates into (last read by the gate GEMM of this step)
    if (key_next != nullptr) key_next[(long long)r * KEY_STRIDE] = 0ull;
  }
  __syncthreads();
  trace_end(trace, trace_slot);
}

cudaError_t launch_finalize_greedy(echr_batch* b, int step, int32_t* tokens, float* logp,
                                   cudaStream_t st, bool keys) {
  const int rows = b->n_events;
  if (rows == 0) return cudaSuccess;
  const echr_model* m = b->model;
  const int n_tiles = logit_groups_for(m, b->n_events);
  cudaError_t e = launch_pdl(finalize_greedy_kernel, dim3((rows + FIN_WARPS - 1) / FIN_WARPS), dim3(FIN_WARPS * 32), 0,
                             st, rows, n_tiles, step, m->cfg.seq_length, b->partial, b->row_tok,
                             b->unfinished, tokens, logp, keys ? b->argmax_key[(step + 1) & 1] : nullptr,
                             trace_ptr(b), trace_next(b, 4));
  count_launch();
  return e;
}

// Teacher-forced step (OldModel_NEW.py:98-130): the logit GEMM stored raw logits into
// logprobs[r, step, :]; subtract the row's log-sum-exp in place and embed seq[r, step+1].
__global__ void __launch_bounds__(FIN_WARPS * 32)
finalize_tf_kernel(int rows, int n_tiles, int step, int seq_cols, int V1, const float* __restrict__ partial,
                   const int32_t* __restrict__ seq, int32_t* __restrict__ row_tok,
                   float* __restrict__ logprobs) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * FIN_WARPS + warp;
  if (r >= rows) return;
  if (step >= 0) {  // step == -1: only embed seq[:, 0]
    const RowStat s = merge_row_stats<1>(partial + (long long)r * n_tiles * LogitRec<1>::WORDS, n_tiles, lane);
    const float lse = s.max_val + s.log_sum;
    float* row = logprobs + ((long long)r * (seq_cols - 1) + step) * V1;
    for (int c = lane; c < V1; c += 32) row[c] -= lse;
  }
  if (step + 1 < seq_cols - 1) {
    int tok = seq[(long long)r * seq_cols + step + 1];
    tok = min(max(tok, 0), V1 - 1);
    if (lane == 0) row_tok[r] = tok;
  }
}

cudaError_t launch_finalize_tf(const echr_batch* b, int step, const int32_t* seq, int seq_cols,
                               float* logprobs, cudaStream_t st) {
  const int rows = b->n_events;
  if (rows == 0) return cudaSuccess;
  const echr_model* m = b->model;
  const int n_tiles = logit_groups_for(m, b->n_events);
  finalize_tf_kernel<<<(rows + FIN_WARPS - 1) / FIN_WARPS, FIN_WARPS * 32, 0, st>>>(
      rows, n_tiles, step, seq_cols, m->V1, b->partial, seq, b->row_tok, logprobs);
  count_launch();
  return cudaGetLastError();
}

// Teacher-forced step of the loss path: the logit GEMM stored raw logits into the scratch rows `logits`;
// keep only what LanguageModelCriterion gathers (misc/utils.py:62-75): the log-prob of the NEXT label,
// target_logp[r, step] = logits[r, seq[r, step+1]] - logsumexp(logits[r, :]), and embed that label.
__global__ void __launch_bounds__(FIN_WARPS * 32)
finalize_nll_kernel(int rows, int n_tiles, int step, int seq_cols, int V1, const float* __restrict__ partial,
                    const int32_t* __restrict__ seq, int32_t* __restrict__ row_tok,
                    const float* __restrict__ logits, float* __restrict__ target_logp) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * FIN_WARPS + warp;
  if (r >= rows) return;
  int tok = seq[(long long)r * seq_cols + step + 1];
  tok = min(max(tok, 0), V1 - 1);
  if (step >= 0) {  // step == -1: only embed seq[:, 0]
    const RowStat s = merge_row_stats<1>(partial + (long long)r * n_tiles * LogitRec<1>::WORDS, n_tiles, lane);
    if (lane == 0)
      target_logp[(long long)r * (seq_cols - 1) + step] = logits[(long long)r * V1 + tok] - (s.max_val + s.log_sum);
  }
  if (step + 1 < seq_cols - 1 && lane == 0) row_tok[r] = tok;
}

cudaError_t launch_finalize_nll(const echr_batch* b, int step, const int32_t* seq, int seq_cols,
                                const float* logits, float* target_logp, cudaStream_t st) {
  const int rows = b->n_events;
  if (rows == 0) return cudaSuccess;
  const echr_model* m = b->model;
  const int n_tiles = logit_groups_for(m, b->n_events);
  finalize_nll_kernel<<<(rows + FIN_WARPS - 1) / FIN_WARPS, FIN_WARPS * 32, 0, st>>>(
      rows, n_tiles, step, seq_cols, m->V1, b->partial, seq, b->row_tok, logits, target_logp);
  count_launch();
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Explicit-state single step (OldModel.get_logprobs_state, OldModel_NEW.py:133-137): load the
// caller's (h, it) into the activation matrix, and read the new h back out of the planes.
__global__ void __launch_bounds__(128)
load_state_kernel(int rows, int V1, const int32_t* __restrict__ it, const float* __restrict__ h_in,
                  int32_t* __restrict__ row_tok, bf16* __restrict__ act, long long act_plane, int planes) {
  const int r = blockIdx.x;
  bf16* row = act + (long long)r * ACT_LD;
  for (int c = threadIdx.x; c < 3 * HID; c += blockDim.x) {
    const int l = c / HID, u = c - l * HID;
    bf16 h, lo;
    split16(h_in[((long long)l * rows + r) * HID + u], planes == 1, h, lo);
    row[ACT_H + c] = h;
    if (planes > 1) row[act_plane + ACT_H + c] = lo;
  }
  if (threadIdx.x == 0) row_tok[r] = min(max(it[r], 0), V1 - 1);
}

__global__ void __launch_bounds__(128)
export_state_kernel(int rows, const bf16* __restrict__ act, long long act_plane, int planes,
                    float* __restrict__ h_out) {
  const int r = blockIdx.x;
  const bf16* row = act + (long long)r * ACT_LD;
  for (int c = threadIdx.x; c < 3 * HID; c += blockDim.x) {
    const int l = c / HID, u = c - l * HID;
    float v = planes > 1 ? __bfloat162float(row[ACT_H + c]) + __bfloat162float(row[act_plane + ACT_H + c])
                         : f16_value(row[ACT_H + c]);
    h_out[((long long)l * rows + r) * HID + u] = v;
  }
}

cudaError_t launch_load_state(const echr_batch* b, int rows, const int32_t* it, const float* h_in, cudaStream_t st) {
  const echr_model* m = b->model;
  load_state_kernel<<<rows, 128, 0, st>>>(rows, m->V1, it, h_in, b->row_tok, b->act[0].p,
                                          b->act[0].plane_stride(), (int)b->act[0].planes);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_export_state(const echr_batch* b, int rows, float* h_out, cudaStream_t st) {
  export_state_kernel<<<rows, 128, 0, st>>>(rows, b->act[1].p, b->act[1].plane_stride(), (int)b->act[1].planes, h_out);
  count_launch();
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Beam step (spec: oracle/echr_oracle.py beam_search).  One CTA per event, warp b = slot b.
constexpr int BEAM_MAX = 4;
constexpr int BEAM_SEG = 3 * HID;  // 1536 elements: h of the three streams / c of the three streams

__global__ void __launch_bounds__(BEAM_MAX * 32)
finalize_beam_kernel(int beam, int n_tiles, int step, int seq_len, const float* __restrict__ partial,
                     int32_t* __restrict__ row_tok, bf16* __restrict__ act_next,
                     long long act_plane, int planes, int32_t* __restrict__ row_parent,
                     float* __restrict__ beam_score, int32_t* __restrict__ beam_live,
                     const int32_t* __restrict__ hist_tok_in, int32_t* __restrict__ hist_tok_out,
                     const float* __restrict__ hist_lp_in, float* __restrict__ hist_lp_out,
                     float* __restrict__ best_score, int32_t* __restrict__ tokens,
                     float* __restrict__ logp, float* __restrict__ score_out,
                     int32_t* __restrict__ trace_parent, int32_t* __restrict__ trace_tok) {
  constexpr int W = LogitRec<4>::WORDS;
  __shared__ float cand_total[BEAM_MAX][BEAM_MAX];
  __shared__ float cand_lp[BEAM_MAX][BEAM_MAX];
  __shared__ int cand_tok[BEAM_MAX][BEAM_MAX];
  __shared__ int sel_parent[BEAM_MAX], sel_tok[BEAM_MAX], sel_fin[BEAM_MAX];
  __shared__ float sel_lp[BEAM_MAX], sel_total[BEAM_MAX];
  const int e = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row0 = e * beam;

  // ---- 1. per-slot top-`beam` log-probs
  if (warp < beam) {
    const int r = row0 + warp;
    const float* rec = partial + (long long)r * n_tiles * W;
    const RowStat s = merge_row_stats<4>(rec, n_tiles, lane);
    const bool live = beam_live[r] != 0;
    const float base = beam_score[r];
    // k rounds of "best candidate strictly after the previous winner" in (value desc, index asc)
    // order; every lane rescans the records of its groups (L1-resident), so any vocabulary fits.
    float last_v = INFINITY;
    int last_i = -1;
    for (int k = 0; k < beam; ++k) {
      float wv = -INFINITY;
      int wi = 0x7fffffff;
      for (int t = lane; t < n_tiles; t += 32) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const float v = rec[t * W + 1 + u];
          const int i = __float_as_int(rec[t * W + 5 + u]);
          const bool eligible = (v < last_v) || (v == last_v && i > last_i);
          if (eligible && (v > wv || (v == wv && i < wi))) { wv = v; wi = i; }
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, wv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, wi, o);
        if (ov > wv || (ov == wv && oi < wi)) { wv = ov; wi = oi; }
      }
      last_v = wv;
      last_i = wi;
      if (lane == 0) {
        const float lp = (wv - s.max_val) - s.log_sum;
        cand_lp[warp][k] = lp;
        cand_tok[warp][k] = wi;
        cand_total[warp][k] = live ? base + lp : -INFINITY;
      }
    }
  }
  __syncthreads();

  // ---- 2. select the `beam` best proposals (score desc, then parent slot, then rank)
  if (threadIdx.x == 0) {
    bool used[BEAM_MAX][BEAM_MAX];
    for (int b = 0; b < BEAM_MAX; ++b)
      for (int k = 0; k < BEAM_MAX; ++k) used[b][k] = false;
    for (int n = 0; n < beam; ++n) {
      int pb = -1, pk = -1;
      float best = -INFINITY;
      for (int b = 0; b < beam; ++b) {
        if (beam_live[row0 + b] == 0) continue;
        for (int k = 0; k < beam; ++k) {
          if (used[b][k]) continue;
          if (pb < 0 || cand_total[b][k] > best) { best = cand_total[b][k]; pb = b; pk = k; }
        }
      }
      if (pb < 0) {  // no live slot left (only after the last step): keep the slot as it is, dead
        sel_parent[n] = n; sel_tok[n] = 0; sel_lp[n] = 0.f; sel_total[n] = -INFINITY; sel_fin[n] = 1;
        continue;
      }
      used[pb][pk] = true;
      sel_parent[n] = pb;
      sel_tok[n] = cand_tok[pb][pk];
      sel_lp[n] = cand_lp[pb][pk];
      sel_total[n] = best;
      sel_fin[n] = (cand_tok[pb][pk] == 0 || step == seq_len - 1) ? 1 : 0;
    }
  }
  __syncthreads();

  // ---- 3. histories (ping-pong), scores, live flags, result update
  if (warp < beam && lane == 0) {  // the step's decisions, for echr_get_beam_trace (what the parity tests verify)
    trace_parent[row0 + warp] = sel_parent[warp];
    trace_tok[row0 + warp] = sel_tok[warp];
  }
  if (warp < beam) {
    const int r = row0 + warp, pr = row0 + sel_parent[warp];
    for (int t = lane; t < seq_len; t += 32) {
      int tk = 0;
      float lp = 0.f;
      if (t < step) { tk = hist_tok_in[(long long)pr * seq_len + t]; lp = hist_lp_in[(long long)pr * seq_len + t]; }
      else if (t == step) { tk = sel_tok[warp]; lp = sel_lp[warp]; }
      hist_tok_out[(long long)r * seq_len + t] = tk;
      hist_lp_out[(long long)r * seq_len + t] = lp;
    }
  }
  __syncthreads();
  if (warp == 0) {
    float best = best_score[e];
    for (int n = 0; n < beam; ++n) {  // ascending slot: lower slots win ties
      if (sel_fin[n] && sel_total[n] > best) {
        best = sel_total[n];
        for (int t = lane; t < seq_len; t += 32) {
          tokens[(long long)e * seq_len + t] = hist_tok_out[(long long)(row0 + n) * seq_len + t];
          logp[(long long)e * seq_len + t] = hist_lp_out[(long long)(row0 + n) * seq_len + t];
        }
      }
    }
    if (lane == 0) {
      best_score[e] = best;
      score_out[e] = best;
    }
    if (lane < beam) {
      beam_score[row0 + lane] = sel_total[lane];
      beam_live[row0 + lane] = sel_fin[lane] ? 0 : 1;
    }
  }
  if (step + 1 >= seq_len) return;

  // ---- 4. next inputs and state reorder
  if (warp < beam && lane == 0) {
    row_tok[row0 + warp] = sel_tok[warp];
    // the cell state (fp32, two thirds of the state bytes) is not moved: the next gate epilogue reads row r's previous
    // c from row row_parent[r] of the buffer this step wrote (GemmProblem::c_src)
    row_parent[row0 + warp] = row0 + sel_parent[warp];
  }
  bool identity = true;
  for (int n = 0; n < beam; ++n) identity = identity && (sel_parent[n] == n);
  if (identity) return;  // uniform across the CTA
  // Reorder h (the planes of the next activation buffer: a TMA operand has to be in place) by parent slot.  A thread
  // owns a column across all slots of the event (reads the <= 4 values, writes them permuted): no staging, no barriers.
  int par[BEAM_MAX];
#pragma unroll
  for (int n = 0; n < BEAM_MAX; ++n) par[n] = n < beam ? sel_parent[n] : n;
  for (int p = 0; p < planes; ++p) {
    bf16* base = act_next + p * act_plane + (long long)row0 * ACT_LD + ACT_H;
    for (int c = threadIdx.x; c < BEAM_SEG / 8; c += blockDim.x) {  // 16-byte words
      uint4 v[BEAM_MAX];
#pragma unroll
      for (int n = 0; n < BEAM_MAX; ++n)
        if (n < beam) v[n] = *reinterpret_cast<const uint4*>(base + (long long)n * ACT_LD + 8 * c);
#pragma unroll
      for (int n = 0; n < BEAM_MAX; ++n) {
        if (n < beam && par[n] != n) {
          uint4 w = v[0];
#pragma unroll
          for (int q = 1; q < BEAM_MAX; ++q)
            if (par[n] == q) w = v[q];
          *reinterpret_cast<uint4*>(base + (long long)n * ACT_LD + 8 * c) = w;
        }
      }
    }
  }
}

cudaError_t launch_finalize_beam(const echr_batch* b, int beam, int step, int32_t* tokens,
                                 float* logp, float* score, cudaStream_t st) {
  if (b->n_events == 0) return cudaSuccess;
  if (beam < 1 || beam > BEAM_MAX) return cudaErrorInvalidValue;
  const echr_model* m = b->model;
  const int n_tiles = logit_groups_for(m, b->n_events * beam);
  const int S = m->cfg.seq_length;
  const long long rows = (long long)b->n_events * beam;
  const PlaneMat& nxt = b->act[(step + 1) & 1];
  const long long hist = (long long)b->cap.max_rows * S;
  const int in = step & 1, out = in ^ 1;
  finalize_beam_kernel<<<b->n_events, BEAM_MAX * 32, 0, st>>>(
      beam, n_tiles, step, S, b->partial, b->row_tok, nxt.p,
      nxt.plane_stride(), (int)nxt.planes, b->row_parent, b->beam_score,
      b->beam_live, b->beam_hist_tok + in * hist, b->beam_hist_tok + out * hist,
      b->beam_hist_lp + in * hist, b->beam_hist_lp + out * hist, b->best_score, tokens, logp, score,
      b->beam_trace + (long long)step * b->cap.max_rows, b->beam_trace + ((long long)S + step) * b->cap.max_rows);
  count_launch();
  return cudaGetLastError();
}

}  // namespace echr
