// Late-fusion attention step (models/OldModel_NEW.py:376-401, ctx2att hoisted) over 16-bit tables, staged through
// shared memory by the bulk-copy engine — the form BASELINE.json's north_star (2) asks for.
//
//   score_l = alpha . tanh(P_l + q)        P = ctx2att(c3d) + b   fp16 [clips,512], written once per batch
//   w       = softmax_l(score)             over the event's clip window [s0, s0 + L)
//   att     = sum_l w_l * clip_l           clip fp16 [clips,512] (500 used): plane 0 of the ctx2att operand
//
// One WARP owns a decode row at a time.  Its lane 0 keeps a private ring of ATT_SLOTS shared-memory slots filled
// by cp.async.bulk (one slot = the 1 KB projected row + the 1 KB clip row of one clip, two copies on one
// mbarrier); the ring runs ahead across row boundaries (the next row is fetched while the current one is being
// consumed), so a warp always has ATT_SLOTS x 2 KB of HBM/L2 reads in flight and never waits on a load with its
// registers.  The window is consumed in ONE pass with an online softmax (running max / sum / weighted feature
// sum rescaled when the max moves): every byte of the window is read exactly once, no score buffer, no second
// pass over the clip rows.  Algorithmic bytes per event-step: 2048 L + 2048 (q in) + 1000 (att out).
//
// tanh(P + q) = 1 - 2 / (e^{2(P+q)} + 1): one FFMA + EX2 + FADD + RCP + FFMA per element, all in fp32; the
// constant sum_k alpha_k and the alpha_net bias drop out of the softmax.  No clamp is needed: e^{2x} -> inf gives
// rcp -> 0, tanh = 1.  The only 16-bit roundings are those of the stored tables (fp16: 2^-12 relative).
//
// The same device code runs (a) as a stand-alone kernel (attention_f16_kernel) and (b) on four extra warps inside
// the persistent logit GEMM (gemm_tc2.cuh, ATT = true), where the memory-bound attention of step t+1 overlaps the
// tensor-bound vocabulary projection of step t.
#pragma once
#include "ptx.cuh"

namespace echr {

constexpr int ATT_SLOTS = 3;           // clips in flight per warp
constexpr int ATT_SLOT_BYTES = 2048;   // projected row (512 fp16) + clip row (512 fp16)
constexpr int ATT_WARP_SMEM = ATT_SLOTS * ATT_SLOT_BYTES;  // + ATT_SLOTS mbarriers (kept separately)

struct AttF16Params {
  const float* q;                // [rows, 512] fp32: h2att(h1) + b  (NOT exponentiated)
  const __nv_bfloat16* proj;     // [clips, 512] fp16 bits
  const __nv_bfloat16* clip;     // [clips, 512] fp16 bits, columns 500..511 zero
  const float* alpha;            // [512]
  const int32_t* row_start;      // [rows] first clip of the row's window
  const int32_t* row_len;        // [rows] clips in the window (>= 1)
  int rows;
  __nv_bfloat16* out;            // [rows, ld_out] fp16 bits (the ATT block of the activation matrix), may be null
  long long ld_out;
  float* out_f32;                // [rows, 500], may be null
  int* row_counter;              // dynamic row scheduling: zero before the launch; null = static round robin
};

struct AttJob {
  int r, s0, L;
};

// `ring`: shared address of this warp's ATT_SLOTS slots (16-byte aligned); `bars`: shared address of its
// ATT_SLOTS mbarriers (already initialised with count 1, phase 0, and only ever used by this warp);
// warp_id / n_warps: this warp's place in the static round robin (ignored with a row counter).
// `kbase`: number of ring slots this warp has consumed before this call (0 on first use) — returned updated, so
// that a persistent caller can invoke the function several times on the same barriers.
__device__ __forceinline__ unsigned att_f16_warp(const AttF16Params& P, uint32_t ring, uint32_t bars, int warp_id,
                                                 int n_warps, unsigned kbase) {
  const int lane = threadIdx.x & 31;
  constexpr float C2 = 2.0f * 1.4426950408889634f;  // 2 log2(e)
  constexpr float L2E = 1.4426950408889634f;
  int static_next = warp_id;
  auto fetch = [&]() {
    AttJob j;
    int r;
    if (P.row_counter != nullptr) {
      r = 0;
      if (lane == 0) r = atomicAdd(P.row_counter, 1);
      r = __shfl_sync(0xffffffffu, r, 0);
    } else {
      r = static_next;
      static_next += n_warps;
    }
    if (r < P.rows) {
      j.r = r;
      j.s0 = __ldg(P.row_start + r);
      j.L = __ldg(P.row_len + r);
    } else {
      j.r = -1; j.s0 = 0; j.L = 0;
    }
    return j;
  };
  AttJob cur = fetch();
  AttJob nxt = cur.r >= 0 ? fetch() : cur;
  unsigned issued = kbase, consumed = kbase;
  int pj = 0, pl = 0;       // producer cursor: job (0 = cur, 1 = nxt) and clip within it
  bool p_done = false;      // the producer has issued everything of `nxt` as well
  auto issue_one = [&]() -> bool {
    if (p_done) return false;
    const AttJob& jb = pj == 0 ? cur : nxt;
    if (jb.r < 0) { p_done = true; return false; }
    if (lane == 0) {
      const unsigned slot = issued % ATT_SLOTS;
      const uint32_t dst = ring + slot * ATT_SLOT_BYTES;
      const uint32_t bar = bars + 8u * slot;
      fence_proxy_async_smem();  // my lanes' generic reads of this slot (ordered by __syncwarp) before the async write
      mbar_arrive_expect_tx(bar, ATT_SLOT_BYTES);
      const long long row = static_cast<long long>(jb.s0 + pl) * 512;
      bulk_load_1d(dst, P.proj + row, 1024, bar);
      bulk_load_1d(dst + 1024, P.clip + row, 1024, bar);
    }
    ++issued;
    if (++pl == jb.L) {
      pl = 0;
      if (pj == 0) pj = 1; else p_done = true;
    }
    return true;
  };
  // lane's hidden units / clip columns: [8 lane + 256 m, +8), m = 0, 1 (16-byte chunks lane and lane + 32 of a row)
  float av[16];
#pragma unroll
  for (int m = 0; m < 2; ++m) {
    const float4 a0 = __ldg(reinterpret_cast<const float4*>(P.alpha + 256 * m + 8 * lane));
    const float4 a1 = __ldg(reinterpret_cast<const float4*>(P.alpha + 256 * m + 8 * lane + 4));
    av[8 * m + 0] = -2.f * a0.x; av[8 * m + 1] = -2.f * a0.y; av[8 * m + 2] = -2.f * a0.z; av[8 * m + 3] = -2.f * a0.w;
    av[8 * m + 4] = -2.f * a1.x; av[8 * m + 5] = -2.f * a1.y; av[8 * m + 6] = -2.f * a1.z; av[8 * m + 7] = -2.f * a1.w;
  }
  auto unpack8 = [](const uint4& u, float* f) {
    const float2 a = f16x2_value(u.x), b = f16x2_value(u.y), c = f16x2_value(u.z), d = f16x2_value(u.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
  };
  while (cur.r >= 0) {
    float qc[16];
#pragma unroll
    for (int m = 0; m < 2; ++m) {
      const float* qp = P.q + static_cast<long long>(cur.r) * 512 + 256 * m + 8 * lane;
      const float4 q0 = __ldg(reinterpret_cast<const float4*>(qp));
      const float4 q1 = __ldg(reinterpret_cast<const float4*>(qp + 4));
      qc[8 * m + 0] = q0.x * C2; qc[8 * m + 1] = q0.y * C2; qc[8 * m + 2] = q0.z * C2; qc[8 * m + 3] = q0.w * C2;
      qc[8 * m + 4] = q1.x * C2; qc[8 * m + 5] = q1.y * C2; qc[8 * m + 6] = q1.z * C2; qc[8 * m + 7] = q1.w * C2;
    }
    float mx = -INFINITY, sum = 0.f;
    float acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = 0.f;
    for (int l = 0; l < cur.L; ++l) {
      while (issued - consumed < ATT_SLOTS && issue_one()) {}
      const unsigned slot = consumed % ATT_SLOTS;
      mbar_wait(bars + 8u * slot, (consumed / ATT_SLOTS) & 1u);
      const uint32_t sp = ring + slot * ATT_SLOT_BYTES;
      float s = 0.f;
      float cf[16];
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        uint4 pu, cu;
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(pu.x), "=r"(pu.y), "=r"(pu.z), "=r"(pu.w)
                     : "r"(sp + 16u * (lane + 32 * m)));
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(cu.x), "=r"(cu.y), "=r"(cu.z), "=r"(cu.w)
                     : "r"(sp + 1024u + 16u * (lane + 32 * m)));
        float pf[8];
        unpack8(pu, pf);
        unpack8(cu, cf + 8 * m);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float e = exp2f(fmaf(pf[i], C2, qc[8 * m + i]));
          float rc;
          asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(e + 1.0f));
          s = fmaf(av[8 * m + i], rc, s);
        }
      }
      __syncwarp();  // every lane has read the slot: it may be refilled
      ++consumed;
      s = warp_sum(s);
      // online softmax update
      const float m_new = fmaxf(mx, s);
      const float corr = exp2f((mx - m_new) * L2E);  // exp2(-inf) = 0 on the first clip
      const float p = exp2f((s - m_new) * L2E);
      sum = fmaf(sum, corr, p);
#pragma unroll
      for (int i = 0; i < 16; ++i) acc[i] = fmaf(acc[i], corr, p * cf[i]);
      mx = m_new;
    }
    const float inv = sum > 0.f ? 1.0f / sum : 0.f;  // (an empty window gives a zero feature)
#pragma unroll
    for (int m = 0; m < 2; ++m) {
      const int c0 = 256 * m + 8 * lane;
      if (P.out != nullptr) {
        __align__(16) __nv_bfloat16 h[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) h[i] = f16_bits(acc[8 * m + i] * inv);
        *reinterpret_cast<uint4*>(P.out + static_cast<long long>(cur.r) * P.ld_out + c0) = *reinterpret_cast<const uint4*>(h);
      }
      if (P.out_f32 != nullptr) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (c0 + i < 500) P.out_f32[static_cast<long long>(cur.r) * 500 + c0 + i] = acc[8 * m + i] * inv;
      }
    }
    // rotate the job window; the producer may already be inside (or past) `nxt`
    cur = nxt;
    nxt = cur.r >= 0 ? fetch() : cur;
    if (p_done) { p_done = false; pj = 1; pl = 0; }
    else pj = 0;
  }
  return consumed;
}

}  // namespace echr
