// Warp-specialised tcgen05 GEMM for sm_100a:   D[M x N] = A[M x K] * W[N x K]^T   (bf16 in, fp32 acc)
//
//   * A (activations: decode rows) and W (weights, nn.Linear layout [out, in]) are both K-major.
//   * Operands may be given as one or two bf16 "planes" (hi, lo) so that hi+lo carries an fp32
//     value to 16 mantissa bits; the kernel then issues hi*hi (+ lo*hi) (+ hi*lo) into the same
//     TMEM accumulator.  A_TERMS/W_TERMS = 1 is the plain bf16 product.
//   * TMA (SWIZZLE_128B boxes of 64 K-elements) -> smem ring -> tcgen05.mma (M=128, N=BLOCK_N,
//     K=16) issued by one thread -> fp32 accumulator in TMEM -> 4 epilogue warps (tcgen05.ld).
//   * The K loop of one problem may walk over non-contiguous 64-column blocks of a wide
//     activation matrix (table a_kb[]): this is how the LSTM layers read [x_t | context | h].
//   * Up to GEMM_MAX_GROUPS independent problems per launch (blockIdx.z).
//
// Epilogues (one thread owns one output row = TMEM lane, so row reductions need no shuffles):
//   EPI_STORE : + bias, optional tanh, store fp32 and/or bf16 hi/lo planes.
//   EPI_LSTM  : columns are gate-interleaved (i,f,g,o of one hidden unit adjacent); applies the
//               LSTMCell update (reference: nn.LSTMCell in models/OldModel_NEW.py:780-782,
//               809-817) and writes c (fp32) and h (bf16 hi/lo planes).
//   EPI_LOGITS: + bias, mask padded vocabulary, per-row / per-64-column-group {sum-exp, top-K}
//               partials for the fused log-softmax / argmax / beam top-k (reference:
//               OldModel_NEW.py:136,158).
// Eight epilogue warps (two per TMEM lane quarter, each taking half of the tile's columns) keep two
// warps on every scheduler; the bias tile goes through shared memory and operands that do not depend
// on the accumulator (c_prev) are fetched before the wait on the MMAs.
#pragma once
#include <cuda.h>
#include "ptx.cuh"
#include "attention_f16.cuh"

namespace echr {

constexpr int GEMM_BLOCK_M = 128;
constexpr int GEMM_BLOCK_K = 64;
constexpr int GEMM_MAX_KB = 32;
constexpr uint16_t GEMM_KB_HI_ONLY = 0x8000;
constexpr int GEMM_MAX_GROUPS = 3;
constexpr int GEMM_EPI_WARPS = 8;                       // two warps per TMEM lane quarter
constexpr int GEMM_THREADS = 128 + GEMM_EPI_WARPS * 32;  // warps 0-3: TMA, MMA, TMEM alloc, spare
constexpr int LOGIT_GROUP = 64;                          // vocabulary columns per partial record

enum : int { EPI_STORE = 0, EPI_LSTM = 1, EPI_LOGITS = 2 };
enum : int { ACT_NONE = 0, ACT_TANH = 1, ACT_EXP2X = 2, ACT_SIGMOID = 3 };  // ACT_EXP2X: exp(2 * clamp(x, -40, 40))

struct GemmProblem {
  CUtensorMap tmA;  // 3-D {K cols, M rows, planes}, box {64, 128, 1}, SWIZZLE_128B
  CUtensorMap tmW;  // 3-D {K cols, N rows, planes}, box {64, BLOCK_N, 1}, SWIZZLE_128B
  int M;            // valid rows of A / D
  int N;            // valid columns of D (rows of W)
  int num_kb;       // K / 64
  int act;          // EPI_STORE activation
  int epi_store;    // persistent kernel, EPI_LOGITS launch: this problem takes the EPI_STORE epilogue instead
  int ksplit;       // SPLIT == 2 kernels: 2 = the two CTAs of a cluster each take half of the k-blocks of one
                    // tile and exchange partial accumulators through DSMEM; 1 = every CTA owns its tile
  uint16_t a_kb[GEMM_MAX_KB];  // A column-block index for each k-block; bit 15: the block's low-order
                               // plane is identically zero (skip its load and its product)
  const float* bias;           // [N] (EPI_LSTM: gate-interleaved order)
  // EPI_STORE
  float* out_f32;
  long long ld_f32;
  __nv_bfloat16* out_hi;
  __nv_bfloat16* out_lo;
  long long ld_bf;
  // EPI_LSTM
  const float* c_in;   // [M, H]
  const int32_t* c_src;  // optional [M]: row of c_in that output row r continues (beam search: the parent slot's
                         // row, so that the cell state is never physically reordered); null = r itself
  float* c_out;        // [M, H]
  const float* row_bias;  // optional [M, N] per-row gate bias (hoisted constant-context terms); replaces bias
  long long ld_row_bias;
  // optional token-indexed gate table [vocab, N]: W_ih[:, x] . embed[token] (+ bias when there is no
  // row_bias), added to the pre-activations of row r from row tok_table[row_tok[r]].  Replaces bias.
  const int32_t* row_tok;
  const float* tok_table;
  long long ld_tok_table;
  // alternative source of the token: packed arg-max keys written by the logit epilogue of the previous step
  // (see argmax_key below); one key every KEY_STRIDE words
  const unsigned long long* row_key;
  int hidden;          // H (N == 4H)
  // EPI_LOGITS
  float* partial;      // [M, n_groups, REC]   n_groups = ceil(N / BLOCK_N) * (BLOCK_N / LOGIT_GROUP)
  float* logits_out;   // optional [M, ld_logits]
  long long ld_logits;
  // optional (TOPK == 1): running arg-max of the row over all tiles, as one 64-bit atomicMax per thread on
  // argmax_key[row * KEY_STRIDE] = (orderable bits of the value) << 32 | (0xFFFFFFFF - column): the largest value
  // wins, the lowest column among equals (torch.max).  Lets the next step's gate GEMM read its input token
  // without waiting for the finalize kernel.  The key must be zero before the launch.
  unsigned long long* argmax_key;
};

constexpr int KEY_STRIDE = 16;  // one key per 128-byte line: the 90 atomics of a row do not share a line with other rows

__device__ __forceinline__ unsigned long long pack_argmax_key(float v, int col) {
  unsigned int u = __float_as_uint(v);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);  // monotone map float -> unsigned
  return (static_cast<unsigned long long>(u) << 32) | static_cast<unsigned long long>(0xFFFFFFFFu - static_cast<unsigned int>(col));
}
__device__ __forceinline__ int token_of_key(unsigned long long key) {
  return static_cast<int>(0xFFFFFFFFu - static_cast<unsigned int>(key & 0xFFFFFFFFull));
}

struct GemmParams {
  GemmProblem prob[GEMM_MAX_GROUPS];
  unsigned long long* trace;  // optional timeline trace (see ptx.cuh)
  int trace_slot;
  int pdl_late;  // 1: release the programmatic dependents when the accumulator is ready (their launch and
                 // prologue overlap this kernel's epilogue) instead of right after this kernel's prologue
  unsigned long long* trace_detail;  // optional [slot][4]: sum over CTAs of (first stage landed, accumulator
                                     // ready, CTA end) - CTA start in ns, and the CTA count
};

template <int TOPK>
struct LogitRec {
  // words: [0]=sum-exp relative to val[0]; [1..TOPK]=values (descending); [1+TOPK..2*TOPK]=indices
  static constexpr int WORDS = (TOPK == 1) ? 4 : 12;
};

template <int A_TERMS, int W_TERMS, int BLOCK_N>
struct GemmCfg {
  static constexpr int A_PLANE_BYTES = GEMM_BLOCK_M * GEMM_BLOCK_K * 2;  // 16 KB
  static constexpr int W_PLANE_BYTES = BLOCK_N * GEMM_BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_TERMS * A_PLANE_BYTES + W_TERMS * W_PLANE_BYTES;
  static constexpr int STAGES_RAW = (196 * 1024) / STAGE_BYTES;
  static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
  static constexpr int TMEM_COLS = BLOCK_N <= 32 ? 32 : (BLOCK_N <= 64 ? 64 : (BLOCK_N <= 128 ? 128 : 256));
  // stages | barriers (256 B) | bias tile (BLOCK_N floats, 1 KB reserved)
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 + 1024;
};

// Fast forms for the epilogues: one ex2 + one fast reciprocal each, ~1e-6 relative error.  The special-function
// instructions are issued directly in their flush-to-zero form: __expf / __fdividef wrap the same two instructions
// in denormal-range fix-ups (FSETP + two FMUL each: ~10 instructions per value where 4-5 do; a third of the LSTM
// epilogue's instruction stream, though that epilogue turned out not to be issue-bound: 62.8 vs 61.7 us).  Same bits
// over the normal range; e^x below 2^-126 flushes to 0 and 1 / (1 + inf) is 0, which are the limits wanted.
__device__ __forceinline__ float ex2_ftz(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_ftz(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float sigmoid_fast(float x) { return rcp_ftz(1.0f + ex2_ftz(-1.4426950408889634f * x)); }
__device__ __forceinline__ float tanh_fast_acc(float x) {
  return fmaf(-2.0f, rcp_ftz(ex2_ftz(2.8853900817779268f * x) + 1.0f), 1.0f);
}

// store 32 consecutive values of one row as bf16 hi+lo planes, or as one fp16 plane when ol is null
__device__ __forceinline__ void store_planes32(const float* v, __nv_bfloat16* oh, __nv_bfloat16* ol,
                                               int valid) {
  const bool single = ol == nullptr;
  if (valid >= 32 && ((reinterpret_cast<uintptr_t>(oh) & 15) == 0) &&
      (ol == nullptr || (reinterpret_cast<uintptr_t>(ol) & 15) == 0)) {
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      __align__(16) __nv_bfloat16 h[8];
      __align__(16) __nv_bfloat16 l[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) split16(v[8 * g + i], single, h[i], l[i]);
      *reinterpret_cast<uint4*>(oh + 8 * g) = *reinterpret_cast<const uint4*>(h);
      if (ol != nullptr) *reinterpret_cast<uint4*>(ol + 8 * g) = *reinterpret_cast<const uint4*>(l);
    }
  } else {
    for (int i = 0; i < 32; ++i) {
      if (i < valid) {
        __nv_bfloat16 h, l;
        split16(v[i], single, h, l);
        oh[i] = h;
        if (ol != nullptr) ol[i] = l;
      }
    }
  }
}

// SPLIT == 2: launched with cluster dims (2,1,1).  For a problem with ksplit == 2 the CTA pair shares one
// output tile: CTA `rank` runs the k-blocks [rank*num_kb/2, (rank+1)*num_kb/2) and finalises the column half
// [rank*BLOCK_N/2, +BLOCK_N/2) of the tile after receiving the peer's partial sums for that half through
// distributed shared memory.  Halving the bytes a CTA pulls from L2 halves its main loop, which in the
// latency regime (a few hundred rows) is bound by the per-SM L2->SM operand stream.
template <int A_TERMS, int W_TERMS, int BLOCK_N, int EPI, int TOPK, int SPLIT = 1>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ GemmParams P) {
  using Cfg = GemmCfg<A_TERMS, W_TERMS, BLOCK_N>;
  static_assert(SPLIT == 1 || SPLIT == 2, "cluster size 1 or 2");
  static_assert(SPLIT == 1 || EPI != EPI_STORE, "split-K is wired for the LSTM and logit epilogues");
  static_assert(SPLIT == 1 || (BLOCK_N / 2) * 512 <= Cfg::STAGES * Cfg::STAGE_BYTES, "exchange buffer aliases the stage ring");
  const GemmProblem& pb = P.prob[blockIdx.z];
  const bool split = SPLIT == 2 && pb.ksplit == 2;
  const int rank = SPLIT == 2 ? static_cast<int>(blockIdx.x & 1) : 0;  // == %cluster_ctarank for cluster dims (2,1,1)
  const int n_tile = split ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int m_tile = blockIdx.y;
  trace_begin(P.trace, P.trace_slot);
  if (n_tile * BLOCK_N >= pb.N || m_tile * GEMM_BLOCK_M >= pb.M) return;  // whole CTA leaves

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + Cfg::STAGES * Cfg::STAGE_BYTES;
  // barriers: full[s] at +8*s, empty[s] at +64+8*s, tmem_full at +128, tmem ptr at +136; bias at +256
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 64u + 8u * s; };
  const uint32_t tmem_full_bar = bar_base + 128u;
  const uint32_t tmem_ptr_smem = bar_base + 136u;
  const uint32_t peer_done_bar = bar_base + 152u;  // SPLIT: the peer's MMAs have retired (its stage ring is free)
  const uint32_t xfer_bar = bar_base + 160u;       // SPLIT: the peer's partial sums have landed in my exchange buffer
  uint8_t* bar_generic = smem_raw + (bar_base - smem_u32(smem_raw));
  volatile uint32_t* tmem_ptr_generic = reinterpret_cast<volatile uint32_t*>(bar_generic + 136);
  float* sbias = reinterpret_cast<float*>(bar_generic + 256);
  volatile unsigned long long* t_start_smem = reinterpret_cast<volatile unsigned long long*>(bar_generic + 144);
  if (P.trace_detail != nullptr && threadIdx.x == 0) *t_start_smem = global_timer_ns();

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_kb = pb.num_kb;
  const int kb_begin = split ? rank * (num_kb >> 1) : 0;
  const int kb_end = split ? (rank == 0 ? (num_kb >> 1) : num_kb) : num_kb;
  const int col0 = n_tile * BLOCK_N;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&pb.tmA);
    tma_prefetch_desc(&pb.tmW);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Cfg::STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(tmem_full_bar, 1);
    if (SPLIT == 2) {
      mbar_init(peer_done_bar, 1);
      mbar_init(xfer_bar, GEMM_EPI_WARPS * 32);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<Cfg::TMEM_COLS>(tmem_ptr_smem);
  if (warp >= 4) {  // bias tile -> shared memory (hidden behind the main loop)
    const int i = threadIdx.x - 128;
    if (i < BLOCK_N) sbias[i] = (pb.bias != nullptr && col0 + i < pb.N) ? __ldg(pb.bias + col0 + i) : 0.f;
  }
  tc_fence_before_sync();
  __syncthreads();
  if (SPLIT == 2) cluster_sync_all();  // the peer's barriers are initialised before anyone arrives on them
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr_generic;
  // everything above touched only constants (weights, descriptors): let the next kernel start its
  // own prologue, then wait for the previous kernel's results
  if (!P.pdl_late) pdl_trigger();
  pdl_wait();

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait_hot(empty_bar(stage), phase ^ 1u);
        const uint32_t sa = smem_base + stage * Cfg::STAGE_BYTES;
        const uint32_t sw = sa + A_TERMS * Cfg::A_PLANE_BYTES;
        const uint16_t kbe = pb.a_kb[kb];
        const bool hi_only = A_TERMS == 2 && (kbe & GEMM_KB_HI_ONLY) != 0;
        mbar_arrive_expect_tx(full_bar(stage), Cfg::STAGE_BYTES - (hi_only ? Cfg::A_PLANE_BYTES : 0));
        const int a_col = static_cast<int>(kbe & 0x7fff) * GEMM_BLOCK_K;
#pragma unroll
        for (int t = 0; t < A_TERMS; ++t)
          if (t == 0 || !hi_only)
            tma_load_3d(sa + t * Cfg::A_PLANE_BYTES, &pb.tmA, full_bar(stage), a_col,
                        m_tile * GEMM_BLOCK_M, t);
#pragma unroll
        for (int t = 0; t < W_TERMS; ++t)
          tma_load_3d(sw + t * Cfg::W_PLANE_BYTES, &pb.tmW, full_bar(stage), kb * GEMM_BLOCK_K,
                      n_tile * BLOCK_N, t);
        if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (single thread) =====================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_m128(BLOCK_N, A_TERMS > 1);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t accum = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait_hot(full_bar(stage), phase);
        tc_fence_after_sync();
        if (kb == kb_begin && P.trace_detail != nullptr)
          atomicAdd(P.trace_detail + 4 * P.trace_slot, global_timer_ns() - *t_start_smem);
        const uint32_t sa = smem_base + stage * Cfg::STAGE_BYTES;
        const uint32_t sw = sa + A_TERMS * Cfg::A_PLANE_BYTES;
        const uint64_t a_hi = umma_desc_k_sw128(sa);
        const uint64_t w_hi = umma_desc_k_sw128(sw);
        const bool hi_only = A_TERMS == 2 && (pb.a_kb[kb] & GEMM_KB_HI_ONLY) != 0;
#pragma unroll
        for (int k = 0; k < GEMM_BLOCK_K / 16; ++k) {
          const uint64_t koff = static_cast<uint64_t>(k * 2);  // 32 B per K=16 step, >>4
          umma_f16(tmem_base, a_hi + koff, w_hi + koff, idesc, accum);
          accum = 1;
          if (A_TERMS == 2 && !hi_only) {
            const uint64_t a_lo = umma_desc_k_sw128(sa + Cfg::A_PLANE_BYTES);
            umma_f16(tmem_base, a_lo + koff, w_hi + koff, idesc, 1);
          }
          if (W_TERMS == 2) {
            const uint64_t w_lo = umma_desc_k_sw128(sw + Cfg::W_PLANE_BYTES);
            umma_f16(tmem_base, a_hi + koff, w_lo + koff, idesc, 1);
          }
        }
        umma_commit(empty_bar(stage));  // frees the smem slot when these MMAs retire
        if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1u; }
      }
      umma_commit(tmem_full_bar);  // accumulator complete
    }
  } else if (warp >= 4) {
    // ===================== epilogue: 8 warps; thread = one output row x half the columns ========
    // (split tiles: x a quarter of the columns — this CTA finalises only its half of the tile)
    constexpr int HALF = BLOCK_N >= 64 ? BLOCK_N / 2 : BLOCK_N;  // columns per epilogue thread (unsplit)
    constexpr int CH = HALF >= 32 ? 32 : 16;
    constexpr int NCH = HALF / CH;
    const int q = warp & 3;              // TMEM lane quarter this warp may access
    const int half = (warp - 4) >> 2;    // which half of the (CTA's share of the) tile's columns
    const bool active = (BLOCK_N >= 64) || half == 0;
    const int row_in_tile = q * 32 + lane;
    const int row = m_tile * GEMM_BLOCK_M + row_in_tile;
    const bool row_ok = row < pb.M;
    const int cpt = split ? HALF / 2 : HALF;                       // columns this thread finalises
    const int hc0 = split ? rank * HALF + half * cpt : half * HALF;  // first of them (within the tile)
    const int xc0 = half * cpt;          // split: the same column within this CTA's exchange buffer
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const uint32_t t_row = t_lane + hc0;
    // exchange buffer [HALF/4][128 rows] float4 at the start of the (by then idle) stage ring
    const float4* xbuf = reinterpret_cast<const float4*>(smem_raw + (smem_base - smem_u32(smem_raw)));
    auto add_peer = [&](float* v, int c) {  // v[0..31] += the peer's partial sums of chunk c of this thread
      if (SPLIT == 2 && split) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float4 t = xbuf[((xc0 + c * 32) / 4 + i) * GEMM_BLOCK_M + row_in_tile];
          v[4 * i] += t.x; v[4 * i + 1] += t.y; v[4 * i + 2] += t.z; v[4 * i + 3] += t.w;
        }
      }
    };

    // ---- operands that do not depend on the accumulator: fetch before waiting for the MMAs
    float cprev[EPI == EPI_LSTM ? HALF / 4 : 1];
    float rbias[EPI == EPI_LSTM ? HALF : 1];
    bool has_rbias = false;
    if constexpr (EPI == EPI_LSTM) {
      has_rbias = pb.row_bias != nullptr || pb.tok_table != nullptr;
#pragma unroll
      for (int i = 0; i < HALF; ++i) rbias[i] = 0.f;
      if (pb.row_bias != nullptr && row_ok) {
        const float4* rp = reinterpret_cast<const float4*>(
            pb.row_bias + static_cast<long long>(row) * pb.ld_row_bias + col0 + hc0);
#pragma unroll
        for (int i = 0; i < HALF / 4; ++i) {
          if (4 * i < cpt) {
            const float4 t = __ldg(rp + i);
            rbias[4 * i] = t.x; rbias[4 * i + 1] = t.y; rbias[4 * i + 2] = t.z; rbias[4 * i + 3] = t.w;
          }
        }
      }
      if (pb.tok_table != nullptr && row_ok) {
        const int tk = pb.row_key != nullptr ? token_of_key(pb.row_key[static_cast<long long>(row) * KEY_STRIDE])
                                             : __ldg(pb.row_tok + row);
        const float4* tp = reinterpret_cast<const float4*>(
            pb.tok_table + static_cast<long long>(tk) * pb.ld_tok_table + col0 + hc0);
#pragma unroll
        for (int i = 0; i < HALF / 4; ++i) {
          if (4 * i < cpt) {
            const float4 t = __ldg(tp + i);
            rbias[4 * i] += t.x; rbias[4 * i + 1] += t.y; rbias[4 * i + 2] += t.z; rbias[4 * i + 3] += t.w;
          }
        }
      }
      if (row_ok) {
        const float4* cp = reinterpret_cast<const float4*>(
            pb.c_in + static_cast<long long>(pb.c_src != nullptr ? __ldg(pb.c_src + row) : row) * pb.hidden + ((col0 + hc0) >> 2));
#pragma unroll
        for (int i = 0; i < HALF / 16; ++i) {
          if (16 * i < cpt) {
            const float4 c = __ldg(cp + i);
            cprev[4 * i] = c.x; cprev[4 * i + 1] = c.y; cprev[4 * i + 2] = c.z; cprev[4 * i + 3] = c.w;
          }
        }
      } else {
#pragma unroll
        for (int i = 0; i < HALF / 4; ++i) cprev[i] = 0.f;
      }
    }

    mbar_wait(tmem_full_bar, 0);
    tc_fence_after_sync();
    if (P.pdl_late && threadIdx.x == 128) pdl_trigger();
    if (P.trace_detail != nullptr && threadIdx.x == 128)
      atomicAdd(P.trace_detail + 4 * P.trace_slot + 1, global_timer_ns() - *t_start_smem);

    if (SPLIT == 2 && split) {
      // 1. tell the peer that my MMAs have retired (my stage ring may now be overwritten by its partial
      //    sums), wait for the same news from the peer
      const uint32_t peer = static_cast<uint32_t>(rank ^ 1);
      if (threadIdx.x == 128) mbar_arrive_cluster(mapa_shared(peer_done_bar, peer));
      mbar_wait_cluster(peer_done_bar, 0);
      // 2. ship my partial sums of the PEER's column half: thread = row x cpt columns, written as
      //    [column/4][row] float4 so that a warp's stores (and the peer's loads) are conflict-free
      const uint32_t xb_remote = mapa_shared(smem_base, peer);
      const int pc0 = static_cast<int>(peer) * HALF + half * cpt;
#pragma unroll 1
      for (int c = 0; c * 32 < cpt; ++c) {
        float v[32];
        tmem_ld_32x32(t_lane + pc0 + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 8; ++i)
          st_cluster_f32x4(xb_remote + static_cast<uint32_t>((((xc0 + c * 32) / 4 + i) * GEMM_BLOCK_M + row_in_tile) * 16),
                           v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
      }
      mbar_arrive_cluster(mapa_shared(xfer_bar, peer));
      // 3. wait for the peer's partial sums of my half
      mbar_wait_cluster(xfer_bar, 0);
    }

    if (active) {
      if constexpr (EPI == EPI_STORE) {
        const int act = pb.act;
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          float v[32];
          if constexpr (CH == 32) tmem_ld_32x32(t_row + c * CH, v);
          else tmem_ld_32x16(t_row + c * CH, v);
          tmem_ld_wait();
          const int tc = hc0 + c * CH;  // column within the tile
          const int cb = col0 + tc;     // global column
          if (row_ok && cb < pb.N) {
#pragma unroll
            for (int i = 0; i < CH; i += 4) {
              const float4 b4 = *reinterpret_cast<const float4*>(sbias + tc + i);
              v[i] += b4.x; v[i + 1] += b4.y; v[i + 2] += b4.z; v[i + 3] += b4.w;
            }
            if (act == ACT_TANH) {
#pragma unroll
              for (int i = 0; i < CH; ++i) v[i] = tanh_fast_acc(v[i]);
            } else if (act == ACT_EXP2X) {
#pragma unroll
              for (int i = 0; i < CH; ++i) v[i] = __expf(2.0f * fminf(fmaxf(v[i], -40.f), 40.f));
            } else if (act == ACT_SIGMOID) {
#pragma unroll
              for (int i = 0; i < CH; ++i) v[i] = __fdividef(1.0f, 1.0f + __expf(-v[i]));
            }
            const int valid = min(CH, pb.N - cb);
            if (pb.out_f32 != nullptr) {
              float* o = pb.out_f32 + static_cast<long long>(row) * pb.ld_f32 + cb;
              if (valid == CH && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
#pragma unroll
                for (int i = 0; i < CH; i += 4)
                  *reinterpret_cast<float4*>(o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
              } else {
                for (int i = 0; i < CH; ++i)
                  if (i < valid) o[i] = v[i];
              }
            }
            if (pb.out_hi != nullptr) {
              __nv_bfloat16* oh = pb.out_hi + static_cast<long long>(row) * pb.ld_bf + cb;
              __nv_bfloat16* ol =
                  pb.out_lo ? pb.out_lo + static_cast<long long>(row) * pb.ld_bf + cb : nullptr;
              if constexpr (CH == 32) store_planes32(v, oh, ol, valid);
              else {
                for (int i = 0; i < CH; ++i)
                  if (i < valid) {
                    __nv_bfloat16 h, l;
                    split16(v[i], ol == nullptr, h, l);
                    oh[i] = h;
                    if (ol) ol[i] = l;
                  }
              }
            }
          }
        }
      } else if constexpr (EPI == EPI_LSTM) {
        // columns are gate-interleaved: 4 consecutive accumulator columns = i,f,g,o of one unit
        const int H = pb.hidden;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          if (c * 32 >= cpt) break;
          float v[32];
          tmem_ld_32x32(t_row + c * 32, v);
          tmem_ld_wait();
          add_peer(v, c);
          if (row_ok) {
            const int tc = hc0 + c * 32;
            const int j0 = (col0 + tc) >> 2;  // first hidden unit of this chunk (8 units)
            float cn[8];
            __align__(16) __nv_bfloat16 hh[8];
            __align__(16) __nv_bfloat16 hl[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              float4 b = *reinterpret_cast<const float4*>(sbias + tc + 4 * u);
              if (has_rbias) {
                b.x = rbias[32 * c + 4 * u]; b.y = rbias[32 * c + 4 * u + 1];
                b.z = rbias[32 * c + 4 * u + 2]; b.w = rbias[32 * c + 4 * u + 3];
              }
              const float gi = sigmoid_fast(v[4 * u + 0] + b.x);
              const float gf = sigmoid_fast(v[4 * u + 1] + b.y);
              const float gg = tanh_fast_acc(v[4 * u + 2] + b.z);
              const float go = sigmoid_fast(v[4 * u + 3] + b.w);
              const float cc = gf * cprev[8 * c + u] + gi * gg;
              cn[u] = cc;
              split16(go * tanh_fast_acc(cc), pb.out_lo == nullptr, hh[u], hl[u]);
            }
            float4* co = reinterpret_cast<float4*>(pb.c_out + static_cast<long long>(row) * H + j0);
            co[0] = make_float4(cn[0], cn[1], cn[2], cn[3]);
            co[1] = make_float4(cn[4], cn[5], cn[6], cn[7]);
            *reinterpret_cast<uint4*>(pb.out_hi + static_cast<long long>(row) * pb.ld_bf + j0) =
                *reinterpret_cast<const uint4*>(hh);
            if (pb.out_lo != nullptr)
              *reinterpret_cast<uint4*>(pb.out_lo + static_cast<long long>(row) * pb.ld_bf + j0) =
                  *reinterpret_cast<const uint4*>(hl);
          }
        }
      } else {  // EPI_LOGITS: one partial record per (row, 64-column group)
        static_assert(EPI != EPI_LOGITS || (SPLIT == 1 && HALF == LOGIT_GROUP) || (SPLIT == 2 && HALF == 2 * LOGIT_GROUP),
                      "logit records are 64 columns wide (split kernels: every problem must have ksplit == 2)");
        constexpr int NCL = LOGIT_GROUP / 32;
        float tv[TOPK];
        int ti[TOPK];
#pragma unroll
        for (int k = 0; k < TOPK; ++k) { tv[k] = -INFINITY; ti[k] = 0x7fffffff; }
        const int gcol0 = col0 + hc0;
        const bool full = gcol0 + LOGIT_GROUP <= pb.N;  // all columns of this thread are real vocabulary
        // pass 1: bias, mask, optional store, running top-K (ascending column scan: ties keep the
        // lower index, like torch.max on the reference path)
#pragma unroll 1
        for (int c = 0; c < NCL; ++c) {
          float v[32];
          tmem_ld_32x32(t_row + c * 32, v);
          tmem_ld_wait();
          add_peer(v, c);
          const int tc = hc0 + c * 32;
          const int cb = col0 + tc;
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const float4 b4 = *reinterpret_cast<const float4*>(sbias + tc + i);
            v[i] += b4.x; v[i + 1] += b4.y; v[i + 2] += b4.z; v[i + 3] += b4.w;
          }
          if (!full) {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (cb + i >= pb.N) v[i] = -INFINITY;
          }
          if constexpr (TOPK == 1) {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (v[i] > tv[0]) { tv[0] = v[i]; ti[0] = cb + i; }
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const float x = v[i];
              const int col = cb + i;
              if (x > tv[TOPK - 1]) {
#pragma unroll
                for (int k = TOPK - 1; k >= 1; --k) {
                  const bool shift = x > tv[k - 1];
                  const bool here = !shift && (x > tv[k]);
                  tv[k] = shift ? tv[k - 1] : (here ? x : tv[k]);
                  ti[k] = shift ? ti[k - 1] : (here ? col : ti[k]);
                }
                if (x > tv[0]) { tv[0] = x; ti[0] = col; }
              }
            }
          }
          if (row_ok && pb.logits_out != nullptr) {
            float* o = pb.logits_out + static_cast<long long>(row) * pb.ld_logits + cb;
            for (int i = 0; i < 32; ++i)
              if (cb + i < pb.N) o[i] = v[i];
          }
        }
        if constexpr (TOPK == 1) {
          if (pb.argmax_key != nullptr && row_ok && gcol0 < pb.N)
            atomicMax(pb.argmax_key + static_cast<long long>(row) * KEY_STRIDE, pack_argmax_key(tv[0], ti[0]));
        }
        // pass 2: sum of exp relative to the group maximum (exp(-inf) = 0 for masked columns)
        float se = 0.f;
        const float m2 = tv[0] * 1.4426950408889634f;
#pragma unroll 1
        for (int c = 0; c < NCL; ++c) {
          float v[32];
          tmem_ld_32x32(t_row + c * 32, v);
          tmem_ld_wait();
          add_peer(v, c);
          const int tc = hc0 + c * 32;
          const int cb = col0 + tc;
          float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const float4 b4 = *reinterpret_cast<const float4*>(sbias + tc + i);
            const float x0 = v[i] + b4.x, x1 = v[i + 1] + b4.y, x2 = v[i + 2] + b4.z, x3 = v[i + 3] + b4.w;
            const float e0 = exp2f(fmaf(x0, 1.4426950408889634f, -m2));
            const float e1 = exp2f(fmaf(x1, 1.4426950408889634f, -m2));
            const float e2 = exp2f(fmaf(x2, 1.4426950408889634f, -m2));
            const float e3 = exp2f(fmaf(x3, 1.4426950408889634f, -m2));
            s4[0] += (full || cb + i < pb.N) ? e0 : 0.f;
            s4[1] += (full || cb + i + 1 < pb.N) ? e1 : 0.f;
            s4[2] += (full || cb + i + 2 < pb.N) ? e2 : 0.f;
            s4[3] += (full || cb + i + 3 < pb.N) ? e3 : 0.f;
          }
          se += (s4[0] + s4[1]) + (s4[2] + s4[3]);
        }
        if (row_ok) {
          const int groups_per_tile = BLOCK_N / LOGIT_GROUP;
          const int n_groups = ((pb.N + BLOCK_N - 1) / BLOCK_N) * groups_per_tile;
          float* rec = pb.partial + (static_cast<long long>(row) * n_groups +
                                     n_tile * groups_per_tile + hc0 / LOGIT_GROUP) * LogitRec<TOPK>::WORDS;
          if (gcol0 >= pb.N) se = 0.f;  // empty group: neutral record
          rec[0] = se;
#pragma unroll
          for (int k = 0; k < TOPK; ++k) {
            rec[1 + k] = tv[k];
            rec[1 + TOPK + k] = __int_as_float(ti[k]);
          }
        }
      }
    }
    tc_fence_before_sync();
  }

  __syncthreads();
  if (warp == 2) {
    tc_fence_after_sync();
    tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
  }
  trace_end(P.trace, P.trace_slot);
  if (P.trace_detail != nullptr && threadIdx.x == 0) {
    atomicAdd(P.trace_detail + 4 * P.trace_slot + 2, global_timer_ns() - *t_start_smem);
    atomicAdd(P.trace_detail + 4 * P.trace_slot + 3, 1ull);
  }
}

// ---------------------------------------------------------------- host side
// Returns nullptr on success or a static error string.
const char* make_tmap_bf16(CUtensorMap* out, const void* ptr, uint64_t cols, uint64_t rows,
                           uint64_t planes, uint64_t ld_elems, uint64_t plane_stride_elems,
                           uint32_t box_rows);

// Dispatch to the instantiated kernel. grid = (max n_tiles, max m_tiles, groups).
cudaError_t gemm_launch(const GemmParams& P, int groups, int max_m, int max_n, int a_terms,
                        int w_terms, int block_n, int epi, int topk, cudaStream_t stream);
// Cluster (2,1,1) split-K variant: EPI_LSTM with block_n 128 (per-problem ksplit 1 or 2) or EPI_LOGITS with
// block_n 256 (ksplit 2).  grid_x = max over problems of n_tiles * ksplit, rounded up to even.
cudaError_t gemm_launch_split(const GemmParams& P, int groups, int max_m, int grid_x, int a_terms, int w_terms,
                              int block_n, int epi, int topk, cudaStream_t stream);

}  // namespace echr
