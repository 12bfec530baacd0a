// Fused pair pipeline of the event relation encoder (models/MA_attention_8_NEW.py:51-79,108-116):
//
//   position matrix -> position embedding [pairs,512] -> pair_pos_fc1 + tanh -> pair_pos_fc2 -> w[pair,16]
//
// in ONE persistent kernel that never writes a [pairs,512] tensor to HBM (the reference builds the embedding in
// numpy float64 on the host and uploads N*N*2 KB per video; round 1 of this library wrote it and the fc1 hidden
// layer to HBM as GEMM operands).  Per tile of 128 event pairs a CTA
//   1. generates the sin/cos embedding straight into shared memory, 128 of the 512 columns at a time (a ring of
//      "super-chunks": the sine k-block and the cosine k-block of 64 frequencies), in the K-major SWIZZLE_128B
//      layout a tcgen05 A operand wants;
//   2. streams pair_pos_fc1.weight through a TMA ring and accumulates the 128 x 512 hidden tile in TMEM (all 512
//      columns: one CTA per SM);
//   3. epilogue 1: hidden = tanh(acc + b1), 64 columns at a time, as 16-bit operand planes into a second ring;
//   4. each hidden chunk feeds a small tcgen05 product (N = 16) against pair_pos_fc2.weight, resident in shared
//      memory, accumulating into TMEM columns 0..15 (free once chunk 0 has been read);
//   5. epilogue 2: + b2 -> pair_w, stored head-major [16][pairs] so that the relation attention reads it coalesced.
// HBM traffic per pair: 64 B out.  Warp roles: 0 TMA producer, 1 MMA issuer, 2 TMEM allocator, 4-11 workers
// (embedding generator and both epilogues).
//
// Embedding arithmetic.  z = 100 * p / 10000^(k/128) reaches ~4e4, so the reference's float64 evaluation cannot
// be replaced by a float32 sincos of z.  Here z / (2 pi) is formed in float64, reduced to [-1/2, 1/2] exactly
// (t - rint(t)), and only the reduced turn goes through float32 sincospif: |error| < 2e-7 against the float64
// value the reference rounds to float32, for ~4 double-precision operations per element instead of a float64
// sincos.
#include "engine.cuh"
#include "util.cuh"

namespace echr {

constexpr int PM_THREADS = 384;
constexpr int PM_WORKERS = 256;
constexpr int PM_ROWS = 128;
constexpr int PM_KB_BYTES = 128 * 128;   // one k-block (64 columns) of one plane: 128 rows x 128 B

template <int A_TERMS, int W_TERMS>
struct PairCfg {
  static constexpr int NH = (A_TERMS == 2 && W_TERMS == 2) ? 128 : 256;  // fc1 output columns per W tile
  static constexpr int PASSES = 512 / NH;
  static constexpr int A_STAGE = 2 * A_TERMS * PM_KB_BYTES;   // sine + cosine k-block
  static constexpr int A_STAGES = A_TERMS == 1 ? 2 : 1;
  static constexpr int H_STAGE = A_TERMS * PM_KB_BYTES;       // 64 hidden columns
  static constexpr int H_STAGES = (A_TERMS == 2 && W_TERMS == 2) ? 1 : 2;
  static constexpr int W_STAGE = W_TERMS * NH * 128;
  static constexpr int W_STAGES = A_TERMS == 1 ? 3 : 2;
  static constexpr int W2_PLANE = 8 * 16 * 128;
  static constexpr int W2_BYTES = W_TERMS * W2_PLANE;
  static constexpr int AUX_BYTES = 256 /*barriers*/ + 2048 /*b1*/ + 64 /*b2*/ + 2048 /*zs*/;
  static constexpr int SMEM_BYTES = 1024 + A_STAGES * A_STAGE + H_STAGES * H_STAGE + W_STAGES * W_STAGE + W2_BYTES + AUX_BYTES;
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
};

struct PairParams {
  CUtensorMap tmW1;  // pair_pos_fc1 planes [512 out, 512 in], box {64, NH, 1}
  CUtensorMap tmW2;  // pair_pos_fc2 planes [16 out, 512 in],  box {64, 16, 1}
  const float* b1;
  const float* b2;
  const int32_t *ev_start, *ev_len, *ev_pair_off, *ev_first;
  int n_events;
  int n_pairs;
  float* pair_w;  // [16][n_pairs]
};

__constant__ double c_pos_invdiv[128];  // 10000^(-k/128)

// byte offset of the 16-byte chunk holding columns [8*chunk, 8*chunk+8) of row r inside a swizzled k-block
__device__ __forceinline__ uint32_t sw128(int r, int chunk) {
  return static_cast<uint32_t>(r * 128 + ((chunk ^ (r & 7)) << 4));
}

__device__ __forceinline__ void pm_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// pair_pos_fc1 k-block that multiplies k-block j (0 sine, 1 cosine) of embedding super-chunk c:
// columns of a pair row are [sin dc (128) | cos dc (128) | sin dl (128) | cos dl (128)]  (MA:51-64)
__device__ __forceinline__ int pm_w_kb(int c, int j) { return (c >> 1) * 4 + 2 * j + (c & 1); }

template <int A_TERMS, int W_TERMS>
__global__ void __launch_bounds__(PM_THREADS, 1)
pair_mlp_kernel(const __grid_constant__ PairParams P) {
  using Cfg = PairCfg<A_TERMS, W_TERMS>;
  constexpr bool SINGLE = A_TERMS == 1;  // one fp16 plane instead of bf16 hi + lo
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* base_generic = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t sA = smem_base;
  const uint32_t sH = sA + Cfg::A_STAGES * Cfg::A_STAGE;
  const uint32_t sW = sH + Cfg::H_STAGES * Cfg::H_STAGE;
  const uint32_t sW2 = sW + Cfg::W_STAGES * Cfg::W_STAGE;
  const uint32_t sAux = sW2 + Cfg::W2_BYTES;
  uint8_t* aux = base_generic + (sAux - smem_base);
  // barriers (8 B each): w_full[4] 0, w_empty[4] 32, a_full[2] 64, a_empty[2] 80, h_full[2] 96, h_empty[2] 112,
  // acc1 128, acc2 136, w2 144; tmem pointer 152
  auto w_full = [&](int s) { return sAux + 8u * s; };
  auto w_empty = [&](int s) { return sAux + 32u + 8u * s; };
  auto a_full = [&](int s) { return sAux + 64u + 8u * s; };
  auto a_empty = [&](int s) { return sAux + 80u + 8u * s; };
  auto h_full = [&](int s) { return sAux + 96u + 8u * s; };
  auto h_empty = [&](int s) { return sAux + 112u + 8u * s; };
  const uint32_t acc1_bar = sAux + 128u, acc2_bar = sAux + 136u, w2_bar = sAux + 144u;
  const uint32_t tmem_ptr_smem = sAux + 152u;
  volatile uint32_t* tmem_ptr_generic = reinterpret_cast<volatile uint32_t*>(aux + 152);
  float* sb1 = reinterpret_cast<float*>(aux + 256);
  float* sb2 = reinterpret_cast<float*>(aux + 256 + 2048);
  double* zs = reinterpret_cast<double*>(aux + 256 + 2048 + 64);  // [2][128]: 100 * p / (2 pi)
  uint8_t* Ag = base_generic;
  uint8_t* Hg = base_generic + (sH - smem_base);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tiles = (P.n_pairs + PM_ROWS - 1) / PM_ROWS;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&P.tmW1);
    tma_prefetch_desc(&P.tmW2);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Cfg::W_STAGES; ++s) {
      mbar_init(w_full(s), 1);
      mbar_init(w_empty(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(a_full(s), PM_WORKERS);
      mbar_init(a_empty(s), 1);
      mbar_init(h_full(s), PM_WORKERS);
      mbar_init(h_empty(s), 1);
    }
    mbar_init(acc1_bar, 1);
    mbar_init(acc2_bar, 1);
    mbar_init(w2_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_ptr_smem);
  if (warp >= 4) {
    const int wt = threadIdx.x - 128;
    sb1[wt] = __ldg(P.b1 + wt);
    sb1[wt + 256] = __ldg(P.b1 + wt + 256);
    if (wt < ECHR_HEADS) sb2[wt] = __ldg(P.b2 + wt);
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr_generic;

  if (warp == 0) {
    // ===================== TMA producer: fc2 weights once, fc1 weight tiles per (tile, chunk, k-block, pass) ====
    if (lane == 0) {
      mbar_arrive_expect_tx(w2_bar, Cfg::W2_BYTES);
      for (int t = 0; t < W_TERMS; ++t)
        for (int kb = 0; kb < 8; ++kb)
          tma_load_3d(sW2 + t * Cfg::W2_PLANE + kb * 2048, &P.tmW2, w2_bar, kb * 64, 0, t);
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int c = 0; c < 4; ++c)
          for (int j = 0; j < 2; ++j)
            for (int ps = 0; ps < Cfg::PASSES; ++ps) {
              mbar_wait(w_empty(stage), phase ^ 1u);
              mbar_arrive_expect_tx(w_full(stage), Cfg::W_STAGE);
              const uint32_t dst = sW + stage * Cfg::W_STAGE;
              for (int t = 0; t < W_TERMS; ++t)
                tma_load_3d(dst + t * Cfg::NH * 128, &P.tmW1, w_full(stage), pm_w_kb(c, j) * 64, ps * Cfg::NH, t);
              if (++stage == Cfg::W_STAGES) { stage = 0; phase ^= 1u; }
            }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc1 = umma_idesc_m128(Cfg::NH, A_TERMS > 1);
      constexpr uint32_t idesc2 = umma_idesc_m128(16, A_TERMS > 1);
      int wstage = 0, astage = 0, hstage = 0;
      uint32_t wphase = 0, aphase = 0, hphase = 0;
      mbar_wait(w2_bar, 0);
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        // ---- fc1: 4 embedding super-chunks x 2 k-blocks x PASSES weight tiles
        for (int c = 0; c < 4; ++c) {
          mbar_wait(a_full(astage), aphase);
          tc_fence_after_sync();
          for (int j = 0; j < 2; ++j) {
            const uint32_t a_hi_addr = sA + astage * Cfg::A_STAGE + j * A_TERMS * PM_KB_BYTES;
            const uint32_t a_lo_addr = a_hi_addr + PM_KB_BYTES;
            for (int ps = 0; ps < Cfg::PASSES; ++ps) {
              mbar_wait(w_full(wstage), wphase);
              tc_fence_after_sync();
              const uint32_t w_addr = sW + wstage * Cfg::W_STAGE;
              const uint64_t a_hi = umma_desc_k_sw128(a_hi_addr);
              const uint64_t w_hi = umma_desc_k_sw128(w_addr);
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const uint64_t koff = static_cast<uint64_t>(k * 2);
                const uint32_t acc = (c > 0 || j > 0 || k > 0) ? 1u : 0u;
                umma_f16(tmem_base + ps * Cfg::NH, a_hi + koff, w_hi + koff, idesc1, acc);
                if (A_TERMS == 2) umma_f16(tmem_base + ps * Cfg::NH, umma_desc_k_sw128(a_lo_addr) + koff, w_hi + koff, idesc1, 1);
                if (W_TERMS == 2)
                  umma_f16(tmem_base + ps * Cfg::NH, a_hi + koff, umma_desc_k_sw128(w_addr + Cfg::NH * 128) + koff, idesc1, 1);
              }
              umma_commit(w_empty(wstage));
              if (++wstage == Cfg::W_STAGES) { wstage = 0; wphase ^= 1u; }
            }
          }
          umma_commit(a_empty(astage));
          if (++astage == Cfg::A_STAGES) { astage = 0; aphase ^= 1u; }
        }
        umma_commit(acc1_bar);
        // ---- fc2: 8 hidden chunks of 64 columns against the resident pair_pos_fc2 k-blocks
        for (int hc = 0; hc < 8; ++hc) {
          mbar_wait(h_full(hstage), hphase);
          tc_fence_after_sync();
          const uint32_t h_addr = sH + hstage * Cfg::H_STAGE;
          const uint64_t a_hi = umma_desc_k_sw128(h_addr);
          const uint64_t w_hi = umma_desc_k_sw128(sW2 + hc * 2048);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t koff = static_cast<uint64_t>(k * 2);
            umma_f16(tmem_base, a_hi + koff, w_hi + koff, idesc2, (hc > 0 || k > 0) ? 1u : 0u);
            if (A_TERMS == 2) umma_f16(tmem_base, umma_desc_k_sw128(h_addr + PM_KB_BYTES) + koff, w_hi + koff, idesc2, 1);
            if (W_TERMS == 2)
              umma_f16(tmem_base, a_hi + koff, umma_desc_k_sw128(sW2 + Cfg::W2_PLANE + hc * 2048) + koff, idesc2, 1);
          }
          umma_commit(h_empty(hstage));
          if (++hstage == Cfg::H_STAGES) { hstage = 0; hphase ^= 1u; }
        }
        umma_commit(acc2_bar);
      }
    }
  } else if (warp >= 4) {
    // ===================== workers: embedding generator + both epilogues =====================
    const int wt = threadIdx.x - 128;
    const int q = warp & 3;             // TMEM lane quarter of this warp
    const int half = (warp - 4) >> 2;   // which 32 of a hidden chunk's 64 columns this thread converts
    const int row = q * 32 + lane;      // tile row of this thread in the epilogues
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const int kl = wt & 63, rg = wt >> 6;  // generator: frequency within the chunk, row group
    int astage = 0, hstage = 0;
    uint32_t aphase = 0, hphase = 0, tphase = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const long long p0 = static_cast<long long>(tile) * PM_ROWS;
      // ---- per-row position features (MA:66-79): centre distance / length ratio of (query i, key j)
      pm_bar_sync(1, PM_WORKERS);  // zs of the previous tile is no longer read
      if (wt < PM_ROWS) {
        const long long p = p0 + wt;
        double z0 = 0.0, z1 = 0.0;
        if (p < P.n_pairs) {
          int lo = 0, hi = P.n_events - 1;  // last event whose first pair row is <= p
          while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (static_cast<long long>(__ldg(P.ev_pair_off + mid)) <= p) lo = mid; else hi = mid - 1;
          }
          const int i = lo;
          const int ej = __ldg(P.ev_first + i) + static_cast<int>(p - __ldg(P.ev_pair_off + i));
          const int si = __ldg(P.ev_start + i), leni = __ldg(P.ev_len + i);
          const int sj = __ldg(P.ev_start + ej), lenj = __ldg(P.ev_len + ej);
          const double ci = 0.5 * (static_cast<double>(si) + static_cast<double>(si + leni));
          const double cj = 0.5 * (static_cast<double>(sj) + static_cast<double>(sj + lenj));
          const float li = static_cast<float>(leni);
          const double dc = fmax(fabs((ci - cj) / static_cast<double>(li)), 1e-3);
          const float ratio = static_cast<float>(lenj) / li;
          const double dl = static_cast<double>(static_cast<float>(log(static_cast<double>(ratio))));
          const double s = 100.0 * 0.15915494309189535;  // 100 / (2 pi)
          z0 = dc * s;
          z1 = dl * s;
        }
        zs[wt] = z0;
        zs[PM_ROWS + wt] = z1;
      }
      pm_bar_sync(1, PM_WORKERS);
      // ---- embedding super-chunks: thread = frequency kl of the chunk, rows rg, rg+4, ...
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        const double invdiv = c_pos_invdiv[64 * (c & 1) + kl];
        const double* z = zs + (c >> 1) * PM_ROWS;
        mbar_wait(a_empty(astage), aphase ^ 1u);
        uint8_t* st = Ag + astage * Cfg::A_STAGE;
        const uint32_t in_chunk = static_cast<uint32_t>((kl & 7) * 2);
#pragma unroll 4
        for (int r = rg; r < PM_ROWS; r += 4) {
          const double t = z[r] * invdiv;
          const float turn = static_cast<float>(t - rint(t));
          float sn, cs;
          sincospif(2.0f * turn, &sn, &cs);
          __nv_bfloat16 sh, sl, ch, cl;
          split16(sn, SINGLE, sh, sl);
          split16(cs, SINGLE, ch, cl);
          const uint32_t off = sw128(r, kl >> 3) + in_chunk;
          *reinterpret_cast<__nv_bfloat16*>(st + off) = sh;                                  // sine k-block, hi
          *reinterpret_cast<__nv_bfloat16*>(st + A_TERMS * PM_KB_BYTES + off) = ch;          // cosine k-block, hi
          if (A_TERMS == 2) {
            *reinterpret_cast<__nv_bfloat16*>(st + PM_KB_BYTES + off) = sl;
            *reinterpret_cast<__nv_bfloat16*>(st + 3 * PM_KB_BYTES + off) = cl;
          }
        }
        fence_proxy_async_smem();  // generic-proxy stores -> visible to the tensor core's async-proxy reads
        mbar_arrive(a_full(astage));
        if (++astage == Cfg::A_STAGES) { astage = 0; aphase ^= 1u; }
      }
      // ---- epilogue 1: hidden = tanh(acc + b1), 64 columns at a time, as operand planes of the fc2 product
      mbar_wait(acc1_bar, tphase);
      tc_fence_after_sync();
#pragma unroll 1
      for (int hc = 0; hc < 8; ++hc) {
        const int col0 = hc * 64 + half * 32;
        float v[32];
        tmem_ld_32x32(t_lane + col0, v);
        tmem_ld_wait();
        mbar_wait(h_empty(hstage), hphase ^ 1u);
        uint8_t* st = Hg + hstage * Cfg::H_STAGE;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          __align__(16) __nv_bfloat16 h[8];
          __align__(16) __nv_bfloat16 l[8];
#pragma unroll
          for (int i = 0; i < 8; ++i)
            split16(tanh_fast_acc(v[8 * g + i] + sb1[col0 + 8 * g + i]), SINGLE, h[i], l[i]);
          const uint32_t off = sw128(row, 4 * half + g);
          *reinterpret_cast<uint4*>(st + off) = *reinterpret_cast<const uint4*>(h);
          if (A_TERMS == 2) *reinterpret_cast<uint4*>(st + PM_KB_BYTES + off) = *reinterpret_cast<const uint4*>(l);
        }
        fence_proxy_async_smem();
        tc_fence_before_sync();  // my reads of these accumulator columns precede the fc2 products (columns 0..15)
        mbar_arrive(h_full(hstage));
        if (++hstage == Cfg::H_STAGES) { hstage = 0; hphase ^= 1u; }
      }
      // ---- epilogue 2: pair weights, head-major
      mbar_wait(acc2_bar, tphase);
      tc_fence_after_sync();
      if (half == 0) {
        float v[16];
        tmem_ld_32x16(t_lane, v);
        tmem_ld_wait();
        const long long p = p0 + row;
        if (p < P.n_pairs) {
#pragma unroll
          for (int g = 0; g < ECHR_HEADS; ++g) P.pair_w[static_cast<long long>(g) * P.n_pairs + p] = v[g] + sb2[g];
        }
      }
      tc_fence_before_sync();  // ... and precede the next tile's fc1 products (ordered by the a_full arrivals)
      tphase ^= 1u;
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after_sync();
    tmem_dealloc<512>(tmem_base);
  }
}

static cudaError_t init_pair_table() {
  static bool done[MAX_DEVICES] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= MAX_DEVICES) return cudaErrorInvalidDevice;
  if (done[dev]) return cudaSuccess;
  double h[128];
  for (int k = 0; k < 128; ++k) h[k] = 1.0 / pow(10000.0, (4.0 / 512.0) * static_cast<double>(k));
  e = cudaMemcpyToSymbol(c_pos_invdiv, h, sizeof(h));
  if (e == cudaSuccess) done[dev] = true;
  return e;
}

template <int A, int W>
static int launch_pair(const PairParams& P, int grid, cudaStream_t st) {
  using Cfg = PairCfg<A, W>;
  static int attr[MAX_DEVICES] = {};
  ECHR_CUDA(set_max_dynamic_smem(pair_mlp_kernel<A, W>, Cfg::SMEM_BYTES, attr));
  pair_mlp_kernel<A, W><<<grid, PM_THREADS, Cfg::SMEM_BYTES, st>>>(P);
  ECHR_CUDA(cudaGetLastError());
  count_launch();
  return ECHR_OK;
}

int launch_pair_mlp(const echr_batch* b, cudaStream_t st) {
  if (b->n_pairs == 0) return ECHR_OK;
  const echr_model* m = b->model;
  const int at = m->cfg.a_terms, wt = m->cfg.w_terms;
  ECHR_CUDA(init_pair_table());
  PairParams P;
  memset(&P, 0, sizeof(P));
  const int nh = (at == 2 && wt == 2) ? 128 : 256;
  const char* err = make_tmap_bf16(&P.tmW1, m->w_fc1.p, HID, HID, (uint64_t)m->w_fc1.planes, HID,
                                   (uint64_t)m->w_fc1.plane_stride(), (uint32_t)nh);
  if (!err)
    err = make_tmap_bf16(&P.tmW2, m->w_fc2.p, HID, ECHR_HEADS, (uint64_t)m->w_fc2.planes, HID,
                         (uint64_t)m->w_fc2.plane_stride(), ECHR_HEADS);
  if (err) return set_error(ECHR_ERR_CUDA, "pair_mlp: %s", err);
  P.b1 = m->b_fc1;
  P.b2 = m->b_fc2;
  P.ev_start = b->ev_start;
  P.ev_len = b->ev_len;
  P.ev_pair_off = b->ev_pair_off;
  P.ev_first = b->ev_first;
  P.n_events = b->n_events;
  P.n_pairs = b->n_pairs;
  P.pair_w = b->pair_w;
  const int n_tiles = (b->n_pairs + PM_ROWS - 1) / PM_ROWS;
  const int sms = num_sms();
  const int grid = n_tiles < sms ? n_tiles : sms;
  if (at == 1 && wt == 1) return launch_pair<1, 1>(P, grid, st);
  if (at == 2 && wt == 1) return launch_pair<2, 1>(P, grid, st);
  if (at == 2 && wt == 2) return launch_pair<2, 2>(P, grid, st);
  return set_error(ECHR_ERR_INVALID, "pair_mlp: unsupported operand planes (%d, %d)", at, wt);
}

}  // namespace echr
