// Persistent tcgen05 GEMM for large row counts (throughput regime):  D[M x N] = A[M x K] * W[N x K]^T
//
// Same operands, plane products and epilogues as gemm_tc.cuh, different schedule:
//   * 128 x 256 output tiles (UMMA M=128, N=256): per k-block a CTA brings in A 16 KB/plane + W 32 KB/plane
//     for twice the tensor work of the 128 x 128 kernel, i.e. 1/3 fewer L2->SM bytes per FLOP — the
//     per-phase trace of the 128 x 128 kernel showed its main loop running at 0.42 us per k-block
//     against 0.26 us of tensor time, bound by operand traffic.
//   * persistent: one CTA per SM walks a static round-robin list of tiles (all problems of the launch
//     flattened; m fastest so that CTAs running together share the same weight tile).
//   * two TMEM accumulators (2 x 256 columns = all of TMEM): the epilogue of tile i overlaps the main
//     loop of tile i+1 (tmem_full / tmem_empty mbarriers), the TMA ring never drains between tiles.
// Warp roles: 0 TMA producer, 1 MMA issuer, 2 TMEM allocator, 3 L2 prefetcher of the LSTM epilogue's gathers (two-plane
// precisions only), 4..11 epilogue (two warps per TMEM lane quarter; a thread owns one output row x 128 columns).
// The epilogue warps never synchronise with each other: each keeps its own copy of its 128 bias values, and the
// staged LSTM epilogue (single-plane precision) its own shared-memory buffers for the gathered operands.
#pragma once
#include "gemm_tc.cuh"

namespace echr {

constexpr int GEMM2_BLOCK_N = 256;

// LSTM epilogue with its gathered operands staged through shared memory (single-plane precision only: the
// two-plane precisions have no shared memory to spare next to two operand stages).  Per epilogue warp and
// 32-column chunk: 32 rows x 128 B of the token gate table + 32 rows x 32 B of cell state.
constexpr bool gemm2_lstm_staged(int a_terms, int w_terms, int epi) { return epi == EPI_LSTM && a_terms == 1 && w_terms == 1; }
constexpr int GEMM2_BIAS_BYTES = GEMM_EPI_WARPS * (GEMM2_BLOCK_N / 2) * 4;  // a private copy of its 128 bias values per epilogue warp
constexpr int GEMM2_STG_TOK = 32 * 128;
constexpr int GEMM2_STG_C = 32 * 32;
constexpr int GEMM2_STG_WARP = GEMM2_STG_TOK + GEMM2_STG_C;        // 5 KB
constexpr int GEMM2_STG_BYTES = GEMM_EPI_WARPS * GEMM2_STG_WARP;   // 40 KB

template <int A_TERMS, int W_TERMS, bool STAGED = false>
struct Gemm2Cfg {
  static constexpr int A_PLANE_BYTES = GEMM_BLOCK_M * GEMM_BLOCK_K * 2;   // 16 KB
  static constexpr int W_PLANE_BYTES = GEMM2_BLOCK_N * GEMM_BLOCK_K * 2;  // 32 KB
  static constexpr int STAGE_BYTES = A_TERMS * A_PLANE_BYTES + W_TERMS * W_PLANE_BYTES;
  static constexpr int STG_BYTES = STAGED ? GEMM2_STG_BYTES : 0;
  static constexpr int STAGES = (200 * 1024 - STG_BYTES) / STAGE_BYTES;  // (1,1): 4 (3 staged), (2,1): 3, (2,2): 2
  // stages | barriers (256 B) | bias: 128 floats per epilogue warp | epilogue operand staging
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 + GEMM2_BIAS_BYTES + STG_BYTES;
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory");
};

struct Gemm2Tile {
  int g, m_tile, n_tile;
};

// linear tile id -> (problem, m tile, n tile).  N_FASTEST = false: m fastest, CTAs running together share a weight
// tile (vocabulary projection: 29 MB of weights, 12 MB of activations).  N_FASTEST = true: the ~148 tiles in flight
// cover all n tiles of a band of rows, so an activation tile is fetched from HBM once and its other readers hit
// L2 whatever else streams through it (gate GEMMs: the three weight matrices are 11 MB and stay hot, while the
// epilogue's gathers of token rows and cell state — 150 MB per launch — would push a row band out of L2 before
// the next of its 8 n tiles came round).
template <bool N_FASTEST>
__device__ __forceinline__ Gemm2Tile gemm2_tile(const GemmParams& P, int groups, int t) {
  Gemm2Tile r;
  r.g = 0;
  for (; r.g < groups - 1; ++r.g) {
    const int mt = (P.prob[r.g].M + GEMM_BLOCK_M - 1) / GEMM_BLOCK_M;
    const int nt = (P.prob[r.g].N + GEMM2_BLOCK_N - 1) / GEMM2_BLOCK_N;
    if (t < mt * nt) break;
    t -= mt * nt;
  }
  const int mt = (P.prob[r.g].M + GEMM_BLOCK_M - 1) / GEMM_BLOCK_M;
  if constexpr (N_FASTEST) {
    const int nt = (P.prob[r.g].N + GEMM2_BLOCK_N - 1) / GEMM2_BLOCK_N;
    r.m_tile = t / nt;
    r.n_tile = t - r.m_tile * nt;
  } else {
    r.n_tile = t / mt;
    r.m_tile = t - r.n_tile * mt;
  }
  return r;
}

template <int A_TERMS, int W_TERMS, int EPI, int TOPK>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tc2_kernel(const __grid_constant__ GemmParams P, int groups, int total_tiles) {
  constexpr bool STAGED = gemm2_lstm_staged(A_TERMS, W_TERMS, EPI);
  using Cfg = Gemm2Cfg<A_TERMS, W_TERMS, STAGED>;
  constexpr int BN = GEMM2_BLOCK_N;
  trace_begin(P.trace, P.trace_slot);

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + Cfg::STAGES * Cfg::STAGE_BYTES;
  // full[s] +8s | empty[s] +64+8s | tmem_full[b] +128+8b | tmem_empty[b] +144+8b | tmem ptr +160 | bias +256
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 64u + 8u * s; };
  auto tmem_full_bar = [&](int b) { return bar_base + 128u + 8u * b; };
  auto tmem_empty_bar = [&](int b) { return bar_base + 144u + 8u * b; };
  const uint32_t tmem_ptr_smem = bar_base + 160u;
  uint8_t* bar_generic = smem_raw + (bar_base - smem_u32(smem_raw));
  volatile uint32_t* tmem_ptr_generic = reinterpret_cast<volatile uint32_t*>(bar_generic + 160);
  float* sbias = reinterpret_cast<float*>(bar_generic + 256);  // [epilogue warp][BN / 2]
  // +176: index of the tile whose epilogue has started (epilogue thread 0 -> prefetch warp; a monotonic counter, so
  // a prefetch warp that falls behind cannot miss a phase the way a waiter on the tmem_full parities could)
  volatile int* epi_tile = reinterpret_cast<volatile int*>(bar_generic + 176);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int g = 0; g < groups; ++g) {
      tma_prefetch_desc(&P.prob[g].tmA);
      tma_prefetch_desc(&P.prob[g].tmW);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Cfg::STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(tmem_full_bar(b), 1);
      mbar_init(tmem_empty_bar(b), GEMM_EPI_WARPS);
    }
    *epi_tile = -1;
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_ptr_smem);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr_generic;
  // Everything above touched only constants.  When the launch is a programmatic dependent of its stream
  // predecessor (util.cuh: launch_pdl) it may have started early: nothing the predecessor writes (activation
  // rows, cell state, attention queries, arg-max keys) is read before this point.  No explicit trigger: the
  // next kernel is released when this one exits.
  pdl_wait();

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const Gemm2Tile tl = gemm2_tile<EPI == EPI_LSTM>(P, groups, t);
        const GemmProblem& pb = P.prob[tl.g];
        for (int kb = 0; kb < pb.num_kb; ++kb) {
          mbar_wait_hot(empty_bar(stage), phase ^ 1u);
          const uint32_t sa = smem_base + stage * Cfg::STAGE_BYTES;
          const uint32_t sw = sa + A_TERMS * Cfg::A_PLANE_BYTES;
          const uint16_t kbe = pb.a_kb[kb];
          const bool hi_only = A_TERMS == 2 && (kbe & GEMM_KB_HI_ONLY) != 0;
          mbar_arrive_expect_tx(full_bar(stage), Cfg::STAGE_BYTES - (hi_only ? Cfg::A_PLANE_BYTES : 0));
          const int a_col = static_cast<int>(kbe & 0x7fff) * GEMM_BLOCK_K;
#pragma unroll
          for (int p = 0; p < A_TERMS; ++p)
            if (p == 0 || !hi_only)
              tma_load_3d(sa + p * Cfg::A_PLANE_BYTES, &pb.tmA, full_bar(stage), a_col,
                          tl.m_tile * GEMM_BLOCK_M, p);
#pragma unroll
          for (int p = 0; p < W_TERMS; ++p)
            tma_load_3d(sw + p * Cfg::W_PLANE_BYTES, &pb.tmW, full_bar(stage), kb * GEMM_BLOCK_K,
                        tl.n_tile * BN, p);
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (single thread) =====================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_m128(BN, A_TERMS > 1);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++it) {
        const Gemm2Tile tl = gemm2_tile<EPI == EPI_LSTM>(P, groups, t);
        const GemmProblem& pb = P.prob[tl.g];
        const int buf = it & 1;
        mbar_wait(tmem_empty_bar(buf), ((it >> 1) & 1) ^ 1u);  // epilogue drained this accumulator
        tc_fence_after_sync();
        const uint32_t acc = tmem_base + buf * BN;
        uint32_t accum = 0;
        for (int kb = 0; kb < pb.num_kb; ++kb) {
          mbar_wait_hot(full_bar(stage), phase);
          tc_fence_after_sync();
          const uint32_t sa = smem_base + stage * Cfg::STAGE_BYTES;
          const uint32_t sw = sa + A_TERMS * Cfg::A_PLANE_BYTES;
          const uint64_t a_hi = umma_desc_k_sw128(sa);
          const uint64_t w_hi = umma_desc_k_sw128(sw);
          const bool hi_only = A_TERMS == 2 && (pb.a_kb[kb] & GEMM_KB_HI_ONLY) != 0;
#pragma unroll
          for (int k = 0; k < GEMM_BLOCK_K / 16; ++k) {
            const uint64_t koff = static_cast<uint64_t>(k * 2);
            umma_f16(acc, a_hi + koff, w_hi + koff, idesc, accum);
            accum = 1;
            if (A_TERMS == 2 && !hi_only) {
              const uint64_t a_lo = umma_desc_k_sw128(sa + Cfg::A_PLANE_BYTES);
              umma_f16(acc, a_lo + koff, w_hi + koff, idesc, 1);
            }
            if (W_TERMS == 2) {
              const uint64_t w_lo = umma_desc_k_sw128(sw + Cfg::W_PLANE_BYTES);
              umma_f16(acc, a_hi + koff, w_lo + koff, idesc, 1);
            }
          }
          umma_commit(empty_bar(stage));
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1u; }
        }
        umma_commit(tmem_full_bar(buf));
      }
    }
  } else if (warp == 3) {
    // ===================== L2 prefetcher (LSTM epilogue only) =====================
    // The epilogue gathers, per output row and tile, 1 KB of the token gate table (a random 8 KB row of a 141 MB
    // table), 256 B of cell state and, when hoisted, 1 KB of per-row bias.  Inside a decode step none of it is
    // in L2 (the attention kernel streams > 100 MB of tables between two gate launches).  This otherwise idle
    // warp asks L2 for the operands of tile i + 1 when the epilogue of tile i starts — one tile (~10 us) ahead:
    // far enough for HBM, near enough that the lines are still there (asking for all tiles of the CTA up front
    // is 140 MB over the GPU, more than L2 holds: measured slower than no prefetch at all).  Not in the staged
    // form: its asynchronous copies already run a chunk ahead (decode of 4072 rows: 5.27 ms with, 5.28 without).
    if constexpr (EPI == EPI_LSTM && !STAGED) {
      auto prefetch_tile = [&](int t) {
        const Gemm2Tile tl = gemm2_tile<EPI == EPI_LSTM>(P, groups, t);
        const GemmProblem& pb = P.prob[tl.g];
        const int col0 = tl.n_tile * BN;
        const long long row0 = static_cast<long long>(tl.m_tile) * GEMM_BLOCK_M;
        int tk[GEMM_BLOCK_M / 32];  // lane l holds the tokens of rows l, l + 32, l + 64, l + 96 (four loads in flight)
#pragma unroll
        for (int u = 0; u < GEMM_BLOCK_M / 32; ++u) {
          const long long row = min(row0 + lane + 32 * u, static_cast<long long>(pb.M) - 1);
          tk[u] = pb.tok_table == nullptr ? 0
                  : pb.row_key != nullptr ? token_of_key(pb.row_key[row * KEY_STRIDE]) : __ldg(pb.row_tok + row);
        }
        // one 128-byte line per lane and instruction: 8 lines per row of the token table / the per-row bias
        // (BN floats), 2 per row of the cell state (BN / 4 floats)
#pragma unroll
        for (int j = 0; j < GEMM_BLOCK_M / 4; ++j) {
          const int r = 4 * j + (lane >> 3);
          const int tkr = __shfl_sync(0xffffffffu, tk[j >> 3], r & 31);
          if (row0 + r < pb.M) {
            if (pb.tok_table != nullptr)
              prefetch_l2_line(pb.tok_table + static_cast<long long>(tkr) * pb.ld_tok_table + col0 + (lane & 7) * 32);
            if (pb.row_bias != nullptr) prefetch_l2_line(pb.row_bias + (row0 + r) * pb.ld_row_bias + col0 + (lane & 7) * 32);
          }
        }
#pragma unroll
        for (int j = 0; j < GEMM_BLOCK_M / 16; ++j) {
          const int r = 16 * j + (lane >> 1);
          if (row0 + r < pb.M)
            prefetch_l2_line(pb.c_in + (pb.c_src != nullptr ? static_cast<long long>(__ldg(pb.c_src + row0 + r)) : row0 + r) * pb.hidden +
                             (col0 >> 2) + (lane & 1) * 32);
        }
      };
      int t = blockIdx.x;
      if (t < total_tiles) {
        prefetch_tile(t);
        for (int it = 0; t + static_cast<int>(gridDim.x) < total_tiles; t += gridDim.x, ++it) {
          while (*epi_tile < it) __nanosleep(200);  // the epilogue of tile `it` has started
          prefetch_tile(t + gridDim.x);
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    constexpr int HALF = BN / 2;  // 128 columns per thread
    constexpr int NCH = HALF / 32;
    const int q = warp & 3;
    const int half = (warp - 4) >> 2;
    const int hc0 = half * HALF;
    const int et = threadIdx.x - 128;  // 0..255
    int it = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++it) {
      const Gemm2Tile tl = gemm2_tile<EPI == EPI_LSTM>(P, groups, t);
      const GemmProblem& pb = P.prob[tl.g];
      const int buf = it & 1;
      const int col0 = tl.n_tile * BN;
      const int row = tl.m_tile * GEMM_BLOCK_M + q * 32 + lane;
      const bool row_ok = row < pb.M;
      // this warp's own copy of the bias of its 128 columns, indexed by the column within the tile like a whole-tile
      // buffer would be: no barrier between the epilogue warps, each runs at its own pace
      float* sb = sbias + (warp - 4) * (BN / 2) - hc0;
      // LSTM epilogue operands that do not depend on the accumulator (see the chunk pipeline below)
      float bbA[EPI == EPI_LSTM ? 32 : 1], bbB[EPI == EPI_LSTM ? 32 : 1], cpA[EPI == EPI_LSTM ? 8 : 1], cpB[EPI == EPI_LSTM ? 8 : 1];
      const bool lstm_rb = EPI == EPI_LSTM && (STAGED || pb.row_bias != nullptr || pb.tok_table != nullptr);
      const long long rowc = row_ok ? row : (pb.M - 1);  // clamped: the loads below are unconditional
      const float* rb_row = nullptr;
      const float* tk_row = nullptr;
      const float* c_row = nullptr;
      if constexpr (EPI == EPI_LSTM) {
        if (pb.row_bias != nullptr) rb_row = pb.row_bias + rowc * pb.ld_row_bias + col0 + hc0;
        if (pb.tok_table != nullptr)
          tk_row = pb.tok_table +
                   static_cast<long long>(pb.row_key != nullptr ? token_of_key(pb.row_key[rowc * KEY_STRIDE])
                                                                : __ldg(pb.row_tok + rowc)) * pb.ld_tok_table + col0 + hc0;
        c_row = pb.c_in + (pb.c_src != nullptr ? static_cast<long long>(__ldg(pb.c_src + rowc)) : rowc) * pb.hidden + ((col0 + hc0) >> 2);
      }
      // Staged form (single-plane precision): the warp copies chunk c of its 32 rows into its own staging buffer
      // with 16-byte cp.async — 8 lanes per 128 B of a token row, 2 per 32 B of cell state: one memory wavefront
      // per 128-byte line, where the row-per-thread 32-byte loads below take one per 32 bytes and run into the
      // load/store unit (measured inside a decode step at 4072 rows: 76 -> 63 us per launch).  Each thread then
      // reads its own row back (16-byte pieces XOR-swizzled by the row so that a quarter-warp covers all banks).
      // The copy of chunk c + 1 is issued as soon as chunk c has been read out and lands during its cell update.
      const uint32_t stg_tok = bar_base + 256u + GEMM2_BIAS_BYTES + static_cast<uint32_t>(warp - 4) * GEMM2_STG_WARP;
      const uint32_t stg_c = stg_tok + GEMM2_STG_TOK;
      int tk_self = 0;
      int csrc_self = static_cast<int>(rowc);  // row of c_in this thread's row continues (GemmProblem::c_src)
      if constexpr (STAGED) {
        if (pb.tok_table != nullptr)
          tk_self = pb.row_key != nullptr ? token_of_key(pb.row_key[rowc * KEY_STRIDE]) : __ldg(pb.row_tok + rowc);
        if (pb.c_src != nullptr) csrc_self = __ldg(pb.c_src + rowc);
      }
      auto stage_chunk = [&](int c) {
        if constexpr (STAGED) {
          const int piece = lane & 7;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int r = 4 * j + (lane >> 3);
            const int tkr = __shfl_sync(0xffffffffu, tk_self, r);
            if (pb.tok_table != nullptr)
              cp_async16(stg_tok + r * 128 + ((piece ^ (r & 7)) << 4),
                         pb.tok_table + static_cast<long long>(tkr) * pb.ld_tok_table + col0 + hc0 + 32 * c + 4 * piece);
          }
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const int r = 16 * j + (lane >> 1);
            const long long rg = __shfl_sync(0xffffffffu, csrc_self, r);  // (clamped to a valid row with its owner)
            cp_async16(stg_c + r * 32 + (((lane & 1) ^ ((r >> 2) & 1)) << 4),
                       pb.c_in + rg * pb.hidden + ((col0 + hc0) >> 2) + 8 * c + 4 * (lane & 1));
          }
          cp_async_commit();
        }
      };
      auto take_chunk = [&](int c, float (&bb)[32], float (&cp)[8]) {
        if constexpr (STAGED) {
          cp_async_wait_all();
          __syncwarp();
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            float4 x;
            if (pb.tok_table != nullptr) {
              x = lds_f32x4(stg_tok + lane * 128 + ((k ^ (lane & 7)) << 4));
            } else {  // no token table (not a decode step): the plain bias
              x = pb.bias != nullptr ? __ldg(reinterpret_cast<const float4*>(pb.bias + col0 + hc0 + 32 * c + 4 * k))
                                     : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            bb[4 * k] = x.x; bb[4 * k + 1] = x.y; bb[4 * k + 2] = x.z; bb[4 * k + 3] = x.w;
          }
#pragma unroll
          for (int k = 0; k < 2; ++k) {
            const float4 x = lds_f32x4(stg_c + lane * 32 + ((k ^ ((lane >> 2) & 1)) << 4));
            cp[4 * k] = x.x; cp[4 * k + 1] = x.y; cp[4 * k + 2] = x.z; cp[4 * k + 3] = x.w;
          }
          if (rb_row != nullptr) {  // hoisted context next to the persistent kernel: forced from a debug knob only
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              float r8[8];
              ldg_nc_f32x8(rb_row + 32 * c + 8 * i, r8);
#pragma unroll
              for (int k = 0; k < 8; ++k) bb[8 * i + k] += r8[k];
            }
          }
          __syncwarp();  // every lane has read its row: the buffer may be refilled
        }
      };
      auto lstm_load = [&](int c, float (&bb)[32], float (&cp)[8]) {
        if constexpr (EPI == EPI_LSTM && !STAGED) {
          ldg_nc_f32x8(c_row + 8 * c, cp);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float r8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (rb_row != nullptr) ldg_nc_f32x8(rb_row + 32 * c + 8 * i, r8);
            if (tk_row != nullptr) {
              float t8[8];
              ldg_nc_f32x8(tk_row + 32 * c + 8 * i, t8);
#pragma unroll
              for (int k = 0; k < 8; ++k) r8[k] += t8[k];
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) bb[8 * i + k] = r8[k];
          }
        }
      };
      if constexpr (STAGED) stage_chunk(0);
      else if constexpr (EPI == EPI_LSTM) lstm_load(0, bbA, cpA);
      // bias of this warp's 128 columns -> its own shared-memory copy
      // (not in the staged form: its bias rides in the token table — or is read directly when there is none — so the
      //  eight epilogue warps never wait for each other)
      if constexpr (!STAGED) {
        __syncwarp();  // every lane is done with the previous tile's values
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int tc = hc0 + 32 * j + lane;
          sb[tc] = (pb.bias != nullptr && col0 + tc < pb.N) ? __ldg(pb.bias + col0 + tc) : 0.f;
        }
        __syncwarp();
      }
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * BN + hc0;
      mbar_wait(tmem_full_bar(buf), (it >> 1) & 1);
      tc_fence_after_sync();
      if (EPI == EPI_LSTM && !STAGED && et == 0) *epi_tile = it;

      // A problem of a logit launch may ask for the plain store epilogue (GemmProblem::epi_store): the attention
      // query GEMM of the next decode step rides in the logit launch of this one (engine.cu: enqueue_decode).
      bool as_store = EPI == EPI_STORE;
      if constexpr (EPI == EPI_LOGITS) as_store = pb.epi_store != 0;
      if (EPI != EPI_LSTM && as_store) {
        const int act = pb.act;
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          float v[32];
          tmem_ld_32x32(t_row + c * 32, v);
          tmem_ld_wait();
          const int tc = hc0 + c * 32;
          const int cb = col0 + tc;
          if (row_ok && cb < pb.N) {
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              const float4 b4 = *reinterpret_cast<const float4*>(sb + tc + i);
              v[i] += b4.x; v[i + 1] += b4.y; v[i + 2] += b4.z; v[i + 3] += b4.w;
            }
            if (act == ACT_TANH) {
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = tanh_fast_acc(v[i]);
            } else if (act == ACT_EXP2X) {
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = __expf(2.0f * fminf(fmaxf(v[i], -40.f), 40.f));
            } else if (act == ACT_SIGMOID) {
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = __fdividef(1.0f, 1.0f + __expf(-v[i]));
            }
            const int valid = min(32, pb.N - cb);
            if (pb.out_f32 != nullptr) {
              float* o = pb.out_f32 + static_cast<long long>(row) * pb.ld_f32 + cb;
              if (valid == 32 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
#pragma unroll
                for (int i = 0; i < 32; i += 4)
                  *reinterpret_cast<float4*>(o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
              } else {
                for (int i = 0; i < 32; ++i)
                  if (i < valid) o[i] = v[i];
              }
            }
            if (pb.out_hi != nullptr) {
              __nv_bfloat16* oh = pb.out_hi + static_cast<long long>(row) * pb.ld_bf + cb;
              __nv_bfloat16* ol =
                  pb.out_lo ? pb.out_lo + static_cast<long long>(row) * pb.ld_bf + cb : nullptr;
              store_planes32(v, oh, ol, valid);
            }
          }
        }
      } else if constexpr (EPI == EPI_LSTM) {
        // Software-pipelined over the four 32-column chunks of this thread: the gathers of chunk c+1 (token
        // gate table row, per-row bias, c_prev: 18 x 16 B per thread, scattered 8 KB rows in HBM / L2) are in
        // flight while chunk c is read from TMEM and pushed through the cell update.  Chunk 0 was requested
        // before the wait on the accumulator (lstm_prefetch above).
        static_assert(NCH == 4, "the chunk pipeline below is unrolled by hand");
        auto lstm_chunk = [&](int c, float (&bb)[32], float (&cp)[8], float (&nbb)[32], float (&ncp)[8]) {
          if constexpr (STAGED) {
            take_chunk(c, bb, cp);
            if (c + 1 < NCH) stage_chunk(c + 1);
          } else {
            if (c + 1 < NCH) lstm_load(c + 1, nbb, ncp);
          }
          const int tc = hc0 + c * 32;
          const int j0 = (col0 + tc) >> 2;  // first hidden unit of this chunk (8 units)
          float v[32];
          tmem_ld_32x32(t_row + c * 32, v);
          tmem_ld_wait();
          if (row_ok) {
            float cn[8];
            __align__(16) __nv_bfloat16 hh[8];
            __align__(16) __nv_bfloat16 hl[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              float b0 = bb[4 * u], b1 = bb[4 * u + 1], b2 = bb[4 * u + 2], b3 = bb[4 * u + 3];
              if (!lstm_rb) {
                const float4 b4 = *reinterpret_cast<const float4*>(sb + tc + 4 * u);
                b0 = b4.x; b1 = b4.y; b2 = b4.z; b3 = b4.w;
              }
              const float gi = sigmoid_fast(v[4 * u + 0] + b0);
              const float gf = sigmoid_fast(v[4 * u + 1] + b1);
              const float gg = tanh_fast_acc(v[4 * u + 2] + b2);
              const float go = sigmoid_fast(v[4 * u + 3] + b3);
              const float cc = gf * cp[u] + gi * gg;
              cn[u] = cc;
              split16(go * tanh_fast_acc(cc), pb.out_lo == nullptr, hh[u], hl[u]);
            }
            stg_f32x8(pb.c_out + static_cast<long long>(row) * pb.hidden + j0, cn);
            *reinterpret_cast<uint4*>(pb.out_hi + static_cast<long long>(row) * pb.ld_bf + j0) =
                *reinterpret_cast<const uint4*>(hh);
            if (pb.out_lo != nullptr)
              *reinterpret_cast<uint4*>(pb.out_lo + static_cast<long long>(row) * pb.ld_bf + j0) =
                  *reinterpret_cast<const uint4*>(hl);
          }
        };
        if constexpr (STAGED) {  // one register set: the next chunk waits in shared memory
          lstm_chunk(0, bbA, cpA, bbA, cpA);
          lstm_chunk(1, bbA, cpA, bbA, cpA);
          lstm_chunk(2, bbA, cpA, bbA, cpA);
          lstm_chunk(3, bbA, cpA, bbA, cpA);
        } else {
          lstm_chunk(0, bbA, cpA, bbB, cpB);
          lstm_chunk(1, bbB, cpB, bbA, cpA);
          lstm_chunk(2, bbA, cpA, bbB, cpB);
          lstm_chunk(3, bbB, cpB, bbA, cpA);
        }
      } else if constexpr (EPI == EPI_LOGITS) {  // one record per (row, 64-column group); this thread owns two groups
        const int n_groups = ((pb.N + BN - 1) / BN) * (BN / LOGIT_GROUP);
#pragma unroll 1
        for (int gq = 0; gq < HALF / LOGIT_GROUP; ++gq) {
          float tv[TOPK];
          int ti[TOPK];
#pragma unroll
          for (int k = 0; k < TOPK; ++k) { tv[k] = -INFINITY; ti[k] = 0x7fffffff; }
          const int gc0 = hc0 + gq * LOGIT_GROUP;  // first column of the group within the tile
          const int gcol0 = col0 + gc0;
          const bool full = gcol0 + LOGIT_GROUP <= pb.N;
#pragma unroll 1
          for (int c = 0; c < LOGIT_GROUP / 32; ++c) {
            float v[32];
            tmem_ld_32x32(t_row + gq * LOGIT_GROUP + c * 32, v);
            tmem_ld_wait();
            const int tc = gc0 + c * 32;
            const int cb = col0 + tc;
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              const float4 b4 = *reinterpret_cast<const float4*>(sb + tc + i);
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
          float se = 0.f;
          const float m2 = tv[0] * 1.4426950408889634f;
#pragma unroll 1
          for (int c = 0; c < LOGIT_GROUP / 32; ++c) {
            float v[32];
            tmem_ld_32x32(t_row + gq * LOGIT_GROUP + c * 32, v);
            tmem_ld_wait();
            const int tc = gc0 + c * 32;
            const int cb = col0 + tc;
            float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              const float4 b4 = *reinterpret_cast<const float4*>(sb + tc + i);
              const float e0 = exp2f(fmaf(v[i] + b4.x, 1.4426950408889634f, -m2));
              const float e1 = exp2f(fmaf(v[i + 1] + b4.y, 1.4426950408889634f, -m2));
              const float e2 = exp2f(fmaf(v[i + 2] + b4.z, 1.4426950408889634f, -m2));
              const float e3 = exp2f(fmaf(v[i + 3] + b4.w, 1.4426950408889634f, -m2));
              s4[0] += (full || cb + i < pb.N) ? e0 : 0.f;
              s4[1] += (full || cb + i + 1 < pb.N) ? e1 : 0.f;
              s4[2] += (full || cb + i + 2 < pb.N) ? e2 : 0.f;
              s4[3] += (full || cb + i + 3 < pb.N) ? e3 : 0.f;
            }
            se += (s4[0] + s4[1]) + (s4[2] + s4[3]);
          }
          if (row_ok) {
            float* rec = pb.partial + (static_cast<long long>(row) * n_groups + tl.n_tile * (BN / LOGIT_GROUP) +
                                       gc0 / LOGIT_GROUP) * LogitRec<TOPK>::WORDS;
            if (gcol0 >= pb.N) se = 0.f;
            rec[0] = se;
#pragma unroll
            for (int k = 0; k < TOPK; ++k) {
              rec[1 + k] = tv[k];
              rec[1 + TOPK + k] = __int_as_float(ti[k]);
            }
          }
        }
      }
      // this accumulator may be overwritten by the MMAs of tile it+2
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(tmem_empty_bar(buf));
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after_sync();
    tmem_dealloc<512>(tmem_base);
  }
  trace_end(P.trace, P.trace_slot);
}

// records per row the logit epilogue of this kernel writes for a vocabulary of n columns
inline int gemm2_logit_groups(int n) { return ((n + GEMM2_BLOCK_N - 1) / GEMM2_BLOCK_N) * (GEMM2_BLOCK_N / LOGIT_GROUP); }
inline int gemm1_logit_groups(int n) { return ((n + 127) / 128) * (128 / LOGIT_GROUP); }

cudaError_t gemm2_launch(const GemmParams& P, int groups, int a_terms, int w_terms, int epi, int topk,
                         int num_sms, cudaStream_t stream);

}  // namespace echr
