// Engine: weight packing (echr_model_*), batch workspace (echr_batch_*), the encode pipeline and
// the device-resident decode loops, exported through the C-ABI of include/echr_b200.h.
#include "engine.cuh"
#include "gemm_tc2.cuh"
#include "util.cuh"

#include <algorithm>
#include <atomic>
#include <new>
#include <vector>

using namespace echr;

namespace echr {

int trace_next(echr_batch* b, int kind) {
  if (b->trace_on != 1 || b->trace_n >= TRACE_SLOTS) return 0;
  b->trace_kind[b->trace_n] = (unsigned char)kind;
  return b->trace_n++;
}

static std::atomic<long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// ---------------------------------------------------------------- packing kernels
// LSTM weights: dst[4j+g, :] = [ w_ih[g*512+j, 0:512] | w_hh[g*512+j, :] | w_ih[g*512+j, 512:512+ctx]
// (zero padded to ctx_pad) ]  — gate-interleaved rows so that one epilogue thread sees i,f,g,o of a
// hidden unit in adjacent accumulator columns; the per-step operands (x, h) come first so that the
// constant-context block (event / video) can be hoisted out of the step GEMM as a column sub-range.
__global__ void pack_lstm_kernel(const float* __restrict__ w_ih, const float* __restrict__ w_hh, int ctx,
                                 int ctx_pad, bf16* __restrict__ hi, bf16* __restrict__ lo, bool f16) {
  const int K = HID + HID + ctx_pad;
  const int in_ih = HID + ctx;
  const long long total = (long long)GATES * K;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i / K), c = (int)(i - (long long)n * K);
    const int j = n >> 2, g = n & 3;
    const int src = g * HID + j;
    float x = 0.f;
    if (c < HID) x = w_ih[(long long)src * in_ih + c];
    else if (c < 2 * HID) x = w_hh[(long long)src * HID + (c - HID)];
    else if (c - 2 * HID < ctx) x = w_ih[(long long)src * in_ih + HID + (c - 2 * HID)];
    bf16 h, l;
    split16(x, f16, h, l);
    hi[i] = h;
    if (lo != nullptr) lo[i] = l;
  }
}
__global__ void pack_lstm_bias_kernel(const float* __restrict__ b_ih, const float* __restrict__ b_hh,
                                      float* __restrict__ out) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n < GATES) {
    const int j = n >> 2, g = n & 3;
    out[n] = b_ih[g * HID + j] + b_hh[g * HID + j];
  }
}
__global__ void copy_f32_kernel(const float* __restrict__ src, float* __restrict__ dst, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src ? src[i] : 0.f;
}

static cudaError_t copy_vec(const float* src, float* dst, int n, cudaStream_t st) {
  copy_f32_kernel<<<(n + 255) / 256, 256, 0, st>>>(src, dst, n);
  return cudaGetLastError();
}

static const char* fill_problem(GemmProblem& pb, const GemmSpec& s, int a_terms, int w_terms, int block_n) {
  if (s.K % GEMM_BLOCK_K != 0 || s.K / GEMM_BLOCK_K > GEMM_MAX_KB) return "bad K";
  if (a_terms > s.A->planes || w_terms > s.W->planes) return "operand has too few planes";
  const char* err = make_tmap_bf16(&pb.tmA, s.A->p, (uint64_t)s.A->ld, (uint64_t)s.M, (uint64_t)s.A->planes,
                                   (uint64_t)s.A->ld, (uint64_t)s.A->plane_stride(), GEMM_BLOCK_M);
  if (err) return err;
  err = make_tmap_bf16(&pb.tmW, s.W->p, (uint64_t)s.W->ld, (uint64_t)s.N, (uint64_t)s.W->planes,
                       (uint64_t)s.W->ld, (uint64_t)s.W->plane_stride(), (uint32_t)block_n);
  if (err) return err;
  pb.M = s.M;
  pb.N = s.N;
  pb.num_kb = s.K / GEMM_BLOCK_K;
  pb.act = s.act;
  pb.epi_store = s.epi_store ? 1 : 0;
  pb.ksplit = s.ksplit;
  for (int i = 0; i < pb.num_kb; ++i)
    pb.a_kb[i] = (uint16_t)((s.a_kb ? s.a_kb[i] : s.a_col0 / GEMM_BLOCK_K + i) | (i < s.hi_only_first ? GEMM_KB_HI_ONLY : 0));
  pb.bias = s.bias;
  pb.out_f32 = s.out_f32;
  pb.ld_f32 = s.ld_f32;
  pb.out_hi = s.out_hi;
  pb.out_lo = s.out_lo;
  pb.ld_bf = s.ld_bf;
  pb.c_in = s.c_in;
  pb.c_src = s.c_src;
  pb.c_out = s.c_out;
  pb.row_bias = s.row_bias;
  pb.ld_row_bias = s.ld_row_bias;
  pb.row_tok = s.row_tok;
  pb.tok_table = s.tok_table;
  pb.ld_tok_table = s.ld_tok_table;
  pb.row_key = s.row_key;
  pb.argmax_key = s.argmax_key;
  pb.hidden = HID;
  pb.partial = s.partial;
  pb.logits_out = s.logits_out;
  pb.ld_logits = s.ld_logits;
  return nullptr;
}

// Row count from which the persistent 128 x 256 kernel (gemm_tc2.cuh) replaces the one-tile-per-CTA
// 128 x 128 kernel: below it there are not enough tiles to fill the SMs with the wide tile.
static int gemm2_min_rows() {
  static int v = -1;
  if (v < 0) {
    const char* e = knob("ECHR_GEMM2_MIN_ROWS");
    v = e ? atoi(e) : 1024;
  }
  return v;
}
static bool use_gemm2(int rows, int n, int block_n) { return block_n == 128 && n >= GEMM2_BLOCK_N && rows >= gemm2_min_rows(); }
// Cluster split-K in the latency regime (ECHR_KSPLIT bit 0: gate GEMMs, bit 1: logit GEMM), used while the
// split launch still fits the SMs in one wave.  Default: gates only — measured on cfg2 (251 rows) the gate
// launch gains 1.3 us per step, while the logit launch (128 x 256 tiles, 64 KB of partial sums per CTA over
// DSMEM at ~17 B/clk) loses 2 us against the unsplit 128 x 128 tiles.
static int ksplit_mask() {
  static int v = -1;
  if (v < 0) {
    const char* e = knob("ECHR_KSPLIT");
    v = e ? atoi(e) : 1;
  }
  return v;
}
// window-length bucket baked into a decode graph (attention variant and its shared-memory staging size)
// (powers of two from 32: few distinct decode graphs for a stream of videos with different spans)
static int len_bucket(int max_len) {
  int bkt = 32;
  while (bkt < max_len) bkt <<= 1;
  return bkt;
}
// Constant-context terms (event / video inputs of streams 0 and 2): hoisted into a per-row gate bias in the
// latency regime, where a gate CTA is bound by its operand stream; kept inside the step GEMM (10 more k-blocks)
// in the throughput regime, where the persistent kernel is bound by the row-per-thread gathers of its epilogue
// and the tensor pipe has room (ECHR_HOIST_CTX=0/1 forces).
static bool hoist_context(int rows) {
  static const int forced = knob("ECHR_HOIST_CTX") ? atoi(knob("ECHR_HOIST_CTX")) : -1;
  if (forced >= 0) return forced != 0;
  return rows < gemm2_min_rows();
}
// throughput regime: the next step's attention query GEMM rides in the logit launch (ECHR_FUSE_Q=0: own launch)
static bool fuse_next_query() {
  static const int v = knob("ECHR_FUSE_Q") ? atoi(knob("ECHR_FUSE_Q")) : 1;
  return v != 0;
}
static int m_tiles_of(int rows) { return (rows + GEMM_BLOCK_M - 1) / GEMM_BLOCK_M; }
// logit GEMM as 128 x 256 tiles shared by CTA pairs (23 x 2 CTAs per 128 rows at the ECHR vocabulary)
static bool logit_split(const echr_model* m, int rows) {
  const int n_tiles = (m->V1 + 255) / 256;
  return (ksplit_mask() & 2) != 0 && rows < gemm2_min_rows() && m_tiles_of(rows) * n_tiles * 2 <= num_sms();
}
// partial records per row the logit GEMM writes for `rows` decode rows
static int logit_groups(const echr_model* m, int rows) {
  return (use_gemm2(rows, m->V1, 128) || logit_split(m, rows)) ? gemm2_logit_groups(m->V1) : gemm1_logit_groups(m->V1);
}

int logit_groups_for(const echr_model* m, int rows) { return logit_groups(m, rows); }

int run_gemm(const GemmSpec* specs, int groups, int a_terms, int w_terms, int block_n, int epi,
             int topk, cudaStream_t st, const char* what, echr_batch* tb, int kind) {
  GemmParams P;
  memset(&P, 0, sizeof(P));
  static const int pdl_late = knob("ECHR_PDL_LATE") ? atoi(knob("ECHR_PDL_LATE")) : 1;
  P.pdl_late = pdl_late;
  if (tb != nullptr && tb->trace_on == 1) { P.trace = tb->trace; P.trace_detail = tb->trace_detail; P.trace_slot = trace_next(tb, kind); }
  int max_m = 0, max_n = 0, min_n = 1 << 30;
  for (int g = 0; g < groups; ++g) {
    if (specs[g].M > max_m) max_m = specs[g].M;
    if (specs[g].N < min_n) min_n = specs[g].N;
  }
  bool any_split = false;
  int grid_x = 0;
  for (int g = 0; g < groups; ++g) {
    any_split = any_split || specs[g].ksplit == 2;
    const int gx = ((specs[g].N + block_n - 1) / block_n) * specs[g].ksplit;
    if (gx > grid_x) grid_x = gx;
  }
  const bool two = !any_split && use_gemm2(max_m, min_n, block_n);
  for (int g = 0; g < groups; ++g)
    if (specs[g].epi_store && !(two && epi == EPI_LOGITS))
      return set_error(ECHR_ERR_INVALID, "%s: a store problem rides only in a persistent logit launch", what);
  for (int g = 0; g < groups; ++g) {
    const char* err = fill_problem(P.prob[g], specs[g], a_terms, w_terms, two ? GEMM2_BLOCK_N : block_n);
    if (err) return set_error(ECHR_ERR_CUDA, "%s: %s", what, err);
    if (specs[g].ksplit == 2 && (P.prob[g].num_kb < 2 || (P.prob[g].num_kb & 1)))
      return set_error(ECHR_ERR_INVALID, "%s: split-K needs an even number of k-blocks", what);
    if (specs[g].M > max_m) max_m = specs[g].M;
    if (specs[g].N > max_n) max_n = specs[g].N;
  }
  if (max_m == 0) return ECHR_OK;
  cudaError_t e = any_split ? gemm_launch_split(P, groups, max_m, grid_x, a_terms, w_terms, block_n, epi, topk, st)
                  : two     ? gemm2_launch(P, groups, a_terms, w_terms, epi, topk, num_sms(), st)
                            : gemm_launch(P, groups, max_m, max_n, a_terms, w_terms, block_n, epi, topk, st);
  if (e != cudaSuccess) return set_error(ECHR_ERR_CUDA, "%s: launch failed: %s", what, cudaGetErrorString(e));
  count_launch();
  return ECHR_OK;
}

int check_device() {
  int dev = 0;
  ECHR_CUDA(cudaGetDevice(&dev));
  return echr_device_check(dev);
}

}  // namespace echr

#define ECHR_TRY(expr)          \
  do {                          \
    int _rc = (expr);           \
    if (_rc != 0) return _rc;   \
  } while (0)

// =============================================================================================
// Model
// =============================================================================================
static void model_layout(echr_model* m, Arena& a) {
  const int wt = m->cfg.w_terms;
  m->w_event_emb = a.planes(wt, HID, FEAT_PAD);
  m->w_qky = a.planes(wt, 3 * HID, HID);
  m->w_fc1 = a.planes(wt, HID, HID);
  m->w_fc2 = a.planes(wt, ECHR_HEADS, HID);
  m->w_ctx2att = a.planes(wt, HID, CLIP_PAD);
  m->w_h2att = a.planes(wt, HID, HID);
  m->w_lstm[0] = a.planes(wt, GATES, HID + HID + HID);       // x | h0 | event
  m->w_lstm[1] = a.planes(wt, GATES, HID + HID + CLIP_PAD);  // x | h1 | attended clip
  m->w_lstm[2] = a.planes(wt, GATES, HID + HID + VID_PAD);   // x | h2 | video
  m->w_logit = a.planes(wt, m->V1, 3 * HID);
  m->embed = a.planes(m->cfg.a_terms == 1 ? 1 : 2, m->V1, HID);  // an A operand of the table build
  m->b_event_emb = a.take<float>(HID);
  m->b_qky = a.take<float>(3 * HID);
  m->b_fc1 = a.take<float>(HID);
  m->b_fc2 = a.take<float>(ECHR_HEADS);
  m->b_ctx2att = a.take<float>(HID);
  m->b_h2att = a.take<float>(HID);
  for (int l = 0; l < 3; ++l) m->b_lstm[l] = a.take<float>(GATES);
  m->b_logit = a.take<float>(m->V1);
  m->b_out = a.take<float>(HID);
  m->alpha = a.take<float>(HID);
  for (int l = 0; l < 3; ++l) m->tok_gates[l] = a.take<float>((size_t)m->V1 * GATES);
}

// f16: the single-plane precision (a_terms == 1) keeps every operand, weights included, as fp16
static cudaError_t pack_linear(const float* w, int out_f, int in_f, const PlaneMat& dst, int row0, bool f16,
                               cudaStream_t st) {
  bf16* hi = dst.plane(0) + (long long)row0 * dst.ld;
  bf16* lo = dst.planes > 1 ? dst.plane(1) + (long long)row0 * dst.ld : nullptr;
  return launch_split(w, out_f, in_f, in_f, hi, lo, dst.ld, dst.ld, f16, st);
}

extern "C" int echr_model_create(const echr_config_t* config, const echr_weights_t* w, void* stream,
                                 echr_model_t** out) {
  ECHR_REQUIRE(config && w && out, "echr_model_create: null argument");
  ECHR_REQUIRE(config->vocab_size >= 1 && config->vocab_size <= 65535, "echr_model_create: vocab_size out of range");
  ECHR_REQUIRE(config->seq_length >= 1 && config->seq_length <= 64, "echr_model_create: seq_length out of range");
  ECHR_REQUIRE((config->a_terms == 1 || config->a_terms == 2) && (config->w_terms == 1 || config->w_terms == 2) &&
                   !(config->a_terms == 1 && config->w_terms == 2),
               "echr_model_create: (a_terms, w_terms) must be (1,1), (2,1) or (2,2)");
  ECHR_REQUIRE(config->clip_bf16 == 0,
               "echr_model_create: clip_bf16 is retired (the single-plane operand format is fp16, whose range cannot "
               "hold the e^{2P} table); pass 0");
  const float* const* wp = reinterpret_cast<const float* const*>(w);
  for (size_t i = 0; i < sizeof(echr_weights_t) / sizeof(float*); ++i)
    ECHR_REQUIRE(wp[i] != nullptr, "echr_model_create: weight pointer %d is null", (int)i);
  ECHR_TRY(check_device());
  cudaStream_t st = (cudaStream_t)stream;
  echr_model* m = new (std::nothrow) echr_model();
  ECHR_REQUIRE(m, "echr_model_create: out of host memory");
  memset(m, 0, sizeof(*m));
  m->cfg = *config;
  m->V1 = config->vocab_size + 1;
  Arena sizing;
  model_layout(m, sizing);
  cudaError_t e = cudaMalloc(&m->arena, sizing.off + 256);
  if (e != cudaSuccess) {
    delete m;
    return set_error(ECHR_ERR_CUDA, "echr_model_create: cudaMalloc(%zu) failed: %s", sizing.off, cudaGetErrorString(e));
  }
  Arena a;
  a.base = reinterpret_cast<char*>(m->arena);
  model_layout(m, a);
#define PK(expr)                                                                        \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess) {                                                            \
      cudaFree(m->arena);                                                               \
      delete m;                                                                         \
      return set_error(ECHR_ERR_CUDA, "echr_model_create: %s: %s", #expr, cudaGetErrorString(_e)); \
    }                                                                                   \
  } while (0)
  PK(cudaMemsetAsync(m->arena, 0, sizing.off, st));
  const bool f16 = config->a_terms == 1;
  PK(pack_linear(w->event_emb_w, HID, ECHR_VIDEO_DIM + HID, m->w_event_emb, 0, f16, st));
  PK(pack_linear(w->query_w, HID, HID, m->w_qky, 0, f16, st));
  PK(pack_linear(w->key_w, HID, HID, m->w_qky, HID, f16, st));
  PK(pack_linear(w->out_w, HID, HID, m->w_qky, 2 * HID, f16, st));
  PK(pack_linear(w->pos_fc1_w, HID, HID, m->w_fc1, 0, f16, st));
  PK(pack_linear(w->pos_fc2_w, ECHR_HEADS, HID, m->w_fc2, 0, f16, st));
  PK(pack_linear(w->ctx2att_w, HID, ECHR_VIDEO_DIM, m->w_ctx2att, 0, f16, st));
  PK(pack_linear(w->h2att_w, HID, HID, m->w_h2att, 0, f16, st));
  PK(pack_linear(w->logit_w, m->V1, 3 * HID, m->w_logit, 0, f16, st));
  PK(pack_linear(w->embed_w, m->V1, HID, m->embed, 0, f16, st));
  const int ctx[3] = {HID, ECHR_VIDEO_DIM, ECHR_LDA_DIM};
  const int ctx_pad[3] = {HID, CLIP_PAD, VID_PAD};
  for (int l = 0; l < 3; ++l) {
    pack_lstm_kernel<<<148 * 8, 256, 0, st>>>(w->lstm_w_ih[l], w->lstm_w_hh[l], ctx[l], ctx_pad[l],
                                              m->w_lstm[l].plane(0),
                                              m->w_lstm[l].planes > 1 ? m->w_lstm[l].plane(1) : nullptr, f16);
    PK(cudaGetLastError());
    pack_lstm_bias_kernel<<<GATES / 256, 256, 0, st>>>(w->lstm_b_ih[l], w->lstm_b_hh[l], m->b_lstm[l]);
    PK(cudaGetLastError());
  }
  PK(copy_vec(w->event_emb_b, m->b_event_emb, HID, st));
  PK(copy_vec(w->query_b, m->b_qky, HID, st));
  PK(copy_vec(w->key_b, m->b_qky + HID, HID, st));
  PK(copy_vec(nullptr, m->b_qky + 2 * HID, HID, st));
  PK(copy_vec(w->pos_fc1_b, m->b_fc1, HID, st));
  PK(copy_vec(w->pos_fc2_b, m->b_fc2, ECHR_HEADS, st));
  PK(copy_vec(w->ctx2att_b, m->b_ctx2att, HID, st));
  PK(copy_vec(w->h2att_b, m->b_h2att, HID, st));
  PK(copy_vec(w->logit_b, m->b_logit, m->V1, st));
  PK(copy_vec(w->out_b, m->b_out, HID, st));
  PK(copy_vec(w->alpha_w, m->alpha, HID, st));
  // token -> gate tables (see engine.cuh): embed [V1,512] (hi+lo planes; one fp16 plane in the single-plane
  // precision) x the x block of each stream's packed weights
  for (int l = 0; l < 3; ++l) {
    PlaneMat wx = m->w_lstm[l];  // columns 0..511 of [x | h | ctx]
    GemmSpec s;
    s.A = &m->embed; s.M = m->V1; s.W = &wx; s.N = GATES; s.K = HID;
    s.bias = m->b_lstm[l];  // b_ih + b_hh ride in the table for every stream
    s.out_f32 = m->tok_gates[l]; s.ld_f32 = GATES;
    const int rc = run_gemm(&s, 1, f16 ? 1 : 2, m->cfg.w_terms, 128, EPI_STORE, 1, st, "token_gate_table");
    if (rc != ECHR_OK) {
      cudaStreamSynchronize(st);
      cudaFree(m->arena);
      delete m;
      return rc;
    }
  }
  PK(cudaStreamSynchronize(st));
#undef PK
  *out = m;
  return ECHR_OK;
}

extern "C" int echr_model_destroy(echr_model_t* m) {
  if (!m) return ECHR_OK;
  cudaFree(m->arena);
  delete m;
  return ECHR_OK;
}

// =============================================================================================
// Batch
// =============================================================================================
static void batch_layout(echr_batch* b, Arena& a) {
  const echr_capacity_t& c = b->cap;
  const echr_model* m = b->model;
  const int ap = m->cfg.a_terms;
  const int S = m->cfg.seq_length;
  const int g1 = gemm1_logit_groups(m->V1), g2 = gemm2_logit_groups(m->V1);
  const int n_tiles = g1 > g2 ? g1 : g2;  // partial records per row (either GEMM variant)
  b->c3d = a.take<float>((size_t)c.max_clips * ECHR_VIDEO_DIM);
  b->tap = a.take<float>((size_t)c.max_events * HID);
  b->lda = a.take<float>((size_t)c.max_videos * ECHR_LDA_DIM);
  int32_t* ints = a.take<int32_t>((size_t)7 * c.max_events + 2 * c.max_videos);
  b->ev_video = ints;
  b->ev_start = ints + 1 * c.max_events;
  b->ev_len = ints + 2 * c.max_events;
  b->ev_ind = ints + 3 * c.max_events;
  b->ev_pair_off = ints + 4 * c.max_events;
  b->ev_first = ints + 5 * c.max_events;
  b->ev_n = ints + 6 * c.max_events;
  b->vid_first = ints + 7 * c.max_events;
  b->vid_n = b->vid_first + c.max_videos;
  if (ap == 1) {  // clip plane = right half of the attention table rows
    b->proj_h = a.take<bf16>((size_t)c.max_clips * 2 * HID);
    b->clip_pl.p = b->proj_h ? b->proj_h + HID : nullptr;
    b->clip_pl.rows = c.max_clips; b->clip_pl.ld = 2 * HID; b->clip_pl.planes = 1; b->clip_pl.stride = 0;
  } else {
    b->proj_h = nullptr;
    b->clip_pl = a.planes(ap, c.max_clips, CLIP_PAD);
  }
  b->proj_f32 = a.take<float>((size_t)c.max_clips * HID);
  b->proj_scale = ap == 1 ? a.take<float>((size_t)c.max_clips) : nullptr;

  b->feat_pl = a.planes(ap, c.max_events, FEAT_PAD);
  b->x_pl = a.planes(ap, c.max_events, HID);
  b->qky = a.take<float>((size_t)c.max_events * 3 * HID);
  b->pair_w = a.take<float>((size_t)c.max_pairs * ECHR_HEADS);
  b->event_ctx = a.take<float>((size_t)c.max_events * HID);
  for (int i = 0; i < 2; ++i) {
    b->act[i] = a.planes(ap, c.max_rows, ACT_LD);
    b->cstate[i] = a.take<float>((size_t)3 * c.max_rows * HID);
  }
  b->q_att = a.take<float>((size_t)c.max_rows * HID);
  for (int i = 0; i < 2; ++i) b->row_bias[i] = a.take<float>((size_t)c.max_rows * GATES);
  b->partial = a.take<float>((size_t)c.max_rows * n_tiles * LogitRec<4>::WORDS);
  b->row_start = a.take<int32_t>(c.max_rows);
  b->row_len = a.take<int32_t>(c.max_rows);
  b->unfinished = a.take<int32_t>(c.max_rows);
  b->row_tok = a.take<int32_t>(c.max_rows);
  b->row_parent = a.take<int32_t>(c.max_rows);
  for (int i = 0; i < 2; ++i) b->argmax_key[i] = a.take<unsigned long long>((size_t)c.max_rows * KEY_STRIDE);
  b->tokens_dev = a.take<int32_t>((size_t)c.max_events * S);
  b->logp_dev = a.take<float>((size_t)c.max_events * S);
  b->score_dev = a.take<float>(c.max_events);
  b->beam_score = a.take<float>(c.max_rows);
  b->beam_live = a.take<int32_t>(c.max_rows);
  b->beam_hist_tok = a.take<int32_t>((size_t)2 * c.max_rows * S);
  b->beam_hist_lp = a.take<float>((size_t)2 * c.max_rows * S);
  b->best_score = a.take<float>(c.max_events);
  b->beam_trace = a.take<int32_t>((size_t)2 * S * c.max_rows);
  b->trace = a.take<unsigned long long>(2 * TRACE_SLOTS);
  b->trace_detail = a.take<unsigned long long>(4 * TRACE_SLOTS);
}

extern "C" int echr_batch_create(const echr_model_t* model, const echr_capacity_t* cap, echr_batch_t** out) {
  ECHR_REQUIRE(model && cap && out, "echr_batch_create: null argument");
  ECHR_REQUIRE(cap->max_videos >= 1 && cap->max_clips >= 1 && cap->max_events >= 1 && cap->max_pairs >= 1 &&
                   cap->max_rows >= cap->max_events,
               "echr_batch_create: capacities must be >= 1 and max_rows >= max_events");
  ECHR_REQUIRE(cap->max_rows < (1ll << 24) && cap->max_pairs < (1ll << 30) && cap->max_clips < (1ll << 30),
               "echr_batch_create: capacity too large");
  ECHR_TRY(check_device());
  echr_batch* b = new (std::nothrow) echr_batch();
  ECHR_REQUIRE(b, "echr_batch_create: out of host memory");
  memset(b, 0, sizeof(*b));
  b->model = model;
  b->cap = *cap;
  Arena sizing;
  batch_layout(b, sizing);
  b->arena_bytes = sizing.off + 256;
  cudaError_t e = cudaMalloc(&b->arena, b->arena_bytes);
  if (e != cudaSuccess) {
    delete b;
    return set_error(ECHR_ERR_CUDA, "echr_batch_create: cudaMalloc(%zu) failed: %s", sizing.off, cudaGetErrorString(e));
  }
  e = cudaMallocHost(reinterpret_cast<void**>(&b->host_stage_base),
                     2 * sizeof(int32_t) * ((size_t)7 * cap->max_events + 2 * cap->max_videos));
  if (e == cudaSuccess)
    e = cudaMallocHost(reinterpret_cast<void**>(&b->host_tap_base), 2 * sizeof(float) * (size_t)cap->max_events * HID);
  for (int i = 0; i < 2 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&b->stage_ev[i], cudaEventDisableTiming);
  if (e != cudaSuccess) {
    if (b->host_stage_base) cudaFreeHost(b->host_stage_base);
    if (b->host_tap_base) cudaFreeHost(b->host_tap_base);
    cudaFree(b->arena);
    delete b;
    return set_error(ECHR_ERR_CUDA, "echr_batch_create: pinned staging allocation failed: %s", cudaGetErrorString(e));
  }
  b->host_stage = b->host_stage_base;
  b->host_tap_stage = b->host_tap_base;
  Arena a;
  a.base = reinterpret_cast<char*>(b->arena);
  batch_layout(b, a);
  e = cudaMemset(b->arena, 0, b->arena_bytes);
  if (e != cudaSuccess) {
    cudaFreeHost(b->host_stage_base);
    cudaFreeHost(b->host_tap_base);
    cudaFree(b->arena);
    delete b;
    return set_error(ECHR_ERR_CUDA, "echr_batch_create: cudaMemset failed: %s", cudaGetErrorString(e));
  }
  *out = b;
  return ECHR_OK;
}

extern "C" int echr_batch_destroy(echr_batch_t* b) {
  if (!b) return ECHR_OK;
  for (int i = 0; i < echr_batch::GRAPH_SLOTS; ++i)
    if (b->graphs[i].exec) cudaGraphExecDestroy(b->graphs[i].exec);
  if (b->cap_stream) {
    cudaStreamDestroy(b->cap_stream);
    cudaStreamDestroy(b->cap_stream2);
    cudaStreamDestroy(b->cap_stream3);
    for (int i = 0; i < 4; ++i) cudaEventDestroy(b->cap_events[i]);
  }
  cudaFreeHost(b->host_stage_base);
  cudaFreeHost(b->host_tap_base);
  for (int i = 0; i < 2; ++i)
    if (b->stage_ev[i]) cudaEventDestroy(b->stage_ev[i]);
  cudaFree(b->nll_logits);
  cudaFree(b->arena);
  delete b;
  return ECHR_OK;
}

// Validates the batch description, derives the per-event / per-video index arrays and uploads them.
// `check_clips` = false is used by the entry points that bring their own (virtual) clip table.
static int stage_desc(echr_batch_t* b, const echr_batch_desc_t* d, bool check_clips, cudaStream_t st) {
  ECHR_REQUIRE(b && d, "echr_batch_set_inputs: null argument");
  ECHR_REQUIRE(d->n_videos >= 1 && d->n_videos <= b->cap.max_videos, "echr_batch_set_inputs: n_videos=%d out of range (capacity %lld)", d->n_videos, (long long)b->cap.max_videos);
  ECHR_REQUIRE(d->n_events >= 0 && d->n_events <= b->cap.max_events, "echr_batch_set_inputs: n_events=%d exceeds capacity %lld", d->n_events, (long long)b->cap.max_events);
  ECHR_REQUIRE(d->host_clip_offsets && (d->n_events == 0 || (d->host_event_video && d->host_event_start && d->host_event_end && d->host_event_ind)),
               "echr_batch_set_inputs: null index array");
  const int32_t* off = d->host_clip_offsets;
  ECHR_REQUIRE(off[0] == 0, "echr_batch_set_inputs: clip_offsets[0] must be 0");
  for (int v = 0; v < d->n_videos; ++v)
    ECHR_REQUIRE(off[v + 1] > off[v], "echr_batch_set_inputs: video %d has no clips", v);
  const int n_clips = off[d->n_videos];
  ECHR_REQUIRE(!check_clips || n_clips <= b->cap.max_clips, "echr_batch_set_inputs: %d clips exceed capacity %lld", n_clips, (long long)b->cap.max_clips);
  // switch to the other half of the pinned staging areas; the copies that read it were enqueued two calls ago
  b->stage_idx ^= 1;
  ECHR_CUDA(cudaEventSynchronize(b->stage_ev[b->stage_idx]));
  const int64_t E = b->cap.max_events;
  b->host_stage = b->host_stage_base + (size_t)b->stage_idx * ((size_t)7 * E + 2 * b->cap.max_videos);
  b->host_tap_stage = b->host_tap_base + (size_t)b->stage_idx * (size_t)E * HID;
  int32_t* hs = b->host_stage;
  int32_t* vid_first = hs + 7 * E;
  int32_t* vid_n = vid_first + b->cap.max_videos;
  for (int v = 0; v < d->n_videos; ++v) { vid_first[v] = 0; vid_n[v] = 0; }
  int prev_v = 0;
  int max_len = 0;
  long long sum_len = 0;
  for (int i = 0; i < d->n_events; ++i) {
    const int v = d->host_event_video[i];
    ECHR_REQUIRE(v >= prev_v && v < d->n_videos, "echr_batch_set_inputs: event %d: video index %d not non-decreasing / out of range", i, v);
    prev_v = v;
    const int T = off[v + 1] - off[v];
    const int s = d->host_event_start[i], e = d->host_event_end[i], ind = d->host_event_ind[i];
    ECHR_REQUIRE(s >= 0 && e > s && e <= T, "echr_batch_set_inputs: event %d: span [%d,%d) invalid for a video of %d clips", i, s, e, T);
    ECHR_REQUIRE(e - s <= ECHR_MAX_SPAN_CLIPS, "echr_batch_set_inputs: event %d: span of %d clips exceeds %d", i, e - s, ECHR_MAX_SPAN_CLIPS);
    ECHR_REQUIRE(ind >= 0 && ind < T, "echr_batch_set_inputs: event %d: ind %d out of range", i, ind);
    if (vid_n[v] == 0) vid_first[v] = i;
    vid_n[v] += 1;
    hs[0 * E + i] = v;
    hs[1 * E + i] = off[v] + s;
    hs[2 * E + i] = e - s;
    if (e - s > max_len) max_len = e - s;
    sum_len += e - s;
    hs[3 * E + i] = off[v] + ind;
  }
  long long pairs = 0;
  int max_n = 0;
  for (int v = 0; v < d->n_videos; ++v) {
    const int n = vid_n[v];
    ECHR_REQUIRE(n <= ECHR_MAX_EVENTS_PER_VIDEO, "echr_batch_set_inputs: video %d has %d events (limit %d)", v, n, ECHR_MAX_EVENTS_PER_VIDEO);
    if (n > max_n) max_n = n;
    for (int j = 0; j < n; ++j) {
      const int i = vid_first[v] + j;
      hs[4 * E + i] = (int32_t)(pairs + (long long)j * n);
      hs[5 * E + i] = vid_first[v];
      hs[6 * E + i] = n;
    }
    pairs += (long long)n * n;
  }
  ECHR_REQUIRE(pairs <= b->cap.max_pairs, "echr_batch_set_inputs: %lld event pairs exceed capacity %lld", pairs, (long long)b->cap.max_pairs);
  b->n_videos = d->n_videos;
  b->n_events = d->n_events;
  b->n_clips = n_clips;
  b->n_pairs = (int)pairs;
  b->max_n = max_n;
  b->max_len = max_len;
  b->mean_len = d->n_events > 0 ? (int)((sum_len + d->n_events - 1) / d->n_events) : 0;
  b->encoded = false;
  ECHR_CUDA(cudaMemcpyAsync(b->ev_video, hs, sizeof(int32_t) * ((size_t)7 * E + 2 * b->cap.max_videos),
                            cudaMemcpyHostToDevice, st));
  ECHR_CUDA(cudaEventRecord(b->stage_ev[b->stage_idx], st));
  return ECHR_OK;
}

extern "C" int echr_batch_set_inputs(echr_batch_t* b, const echr_batch_desc_t* d, const float* c3d,
                                     const float* tap, const float* lda, int on_host, void* stream) {
  ECHR_REQUIRE(b && d && c3d && tap && lda, "echr_batch_set_inputs: null argument");
  cudaStream_t st = (cudaStream_t)stream;
  ECHR_TRY(stage_desc(b, d, true, st));
  const int n_clips = b->n_clips;
  const int64_t E = b->cap.max_events;
  const int32_t* hs = b->host_stage;
  const cudaMemcpyKind kind = on_host ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice;
  ECHR_CUDA(cudaMemcpyAsync(b->c3d, c3d, sizeof(float) * (size_t)n_clips * ECHR_VIDEO_DIM, kind, st));
  // only tap_feats[ind] is ever read (CaptionGenerator.py:121): gather one row per event
  if (on_host) {
    for (int i = 0; i < d->n_events; ++i)
      memcpy(b->host_tap_stage + (size_t)i * HID, tap + (size_t)hs[3 * E + i] * HID, sizeof(float) * HID);
    if (d->n_events > 0)
    {
      ECHR_CUDA(cudaMemcpyAsync(b->tap, b->host_tap_stage, sizeof(float) * (size_t)d->n_events * HID,
                                cudaMemcpyHostToDevice, st));
      ECHR_CUDA(cudaEventRecord(b->stage_ev[b->stage_idx], st));
    }
  } else {
    ECHR_CUDA(launch_gather_tap(b, tap, st));
  }
  ECHR_CUDA(cudaMemcpyAsync(b->lda, lda, sizeof(float) * (size_t)d->n_videos * ECHR_LDA_DIM, kind, st));
  return ECHR_OK;
}

// Graph cache: look `key` up, or capture `enqueue(capture stream)` into a new graph (evicting the least
// recently used slot), then launch it on `st`.
static int ensure_capture_streams(echr_batch* b) {
  if (b->cap_stream != nullptr) return ECHR_OK;
  // capture on private streams: the caller's stream may be the legacy default stream, which cannot be captured
  ECHR_CUDA(cudaStreamCreateWithFlags(&b->cap_stream, cudaStreamNonBlocking));
  ECHR_CUDA(cudaStreamCreateWithFlags(&b->cap_stream2, cudaStreamNonBlocking));
  ECHR_CUDA(cudaStreamCreateWithFlags(&b->cap_stream3, cudaStreamNonBlocking));
  for (int i = 0; i < 4; ++i) ECHR_CUDA(cudaEventCreateWithFlags(&b->cap_events[i], cudaEventDisableTiming));
  return ECHR_OK;
}

// True when `key` has a cached graph or was seen before (second sight: worth capturing); remembers it otherwise.
static bool worth_a_graph(echr_batch* b, const long long (&key)[6]) {
  for (int i = 0; i < echr_batch::GRAPH_SLOTS; ++i)
    if (b->graphs[i].exec != nullptr && memcmp(b->graphs[i].key, key, sizeof(key)) == 0) return true;
  for (int i = 0; i < echr_batch::SEEN_KEYS; ++i)
    if (memcmp(b->seen_keys[i], key, sizeof(key)) == 0) return true;
  memcpy(b->seen_keys[b->seen_next], key, sizeof(key));
  b->seen_next = (b->seen_next + 1) % echr_batch::SEEN_KEYS;
  return false;
}

template <typename F>
static int run_graph(echr_batch* b, const long long (&key)[6], cudaStream_t st, const char* what, bool traced, F&& enqueue) {
  echr_batch::GraphSlot* slot = nullptr;
  for (int i = 0; i < echr_batch::GRAPH_SLOTS; ++i) {
    echr_batch::GraphSlot& g = b->graphs[i];
    if (g.exec != nullptr && memcmp(g.key, key, sizeof(g.key)) == 0) { slot = &g; break; }
  }
  if (slot == nullptr) {
    ECHR_TRY(ensure_capture_streams(b));
    slot = &b->graphs[0];
    for (int i = 0; i < echr_batch::GRAPH_SLOTS; ++i) {
      echr_batch::GraphSlot& g = b->graphs[i];
      if (g.exec == nullptr) { slot = &g; break; }
      if (g.stamp < slot->stamp) slot = &g;
    }
    if (slot->exec) { cudaGraphExecDestroy(slot->exec); slot->exec = nullptr; }
    if (traced) {
      const char* tr = knob("ECHR_TRACE");
      b->trace_on = (tr != nullptr && tr[0] == '1') ? 1 : 0;
      b->trace_n = 0;
      if (b->trace_on)  // the trace slots belong to one graph: drop every other cached graph
        for (int i = 0; i < echr_batch::GRAPH_SLOTS; ++i)
          if (b->graphs[i].exec) { cudaGraphExecDestroy(b->graphs[i].exec); b->graphs[i].exec = nullptr; }
    }
    const int saved_trace = b->trace_on;
    if (!traced) b->trace_on = 0;
    const long long before = g_launches.load();
    cudaGraph_t graph = nullptr;
    ECHR_CUDA(cudaStreamBeginCapture(b->cap_stream, cudaStreamCaptureModeRelaxed));
    const int rc = enqueue(b->cap_stream);
    cudaError_t e = cudaStreamEndCapture(b->cap_stream, &graph);
    if (!traced) b->trace_on = saved_trace;
    if (rc != 0) { if (graph) cudaGraphDestroy(graph); return rc; }
    if (e != cudaSuccess) return set_error(ECHR_ERR_CUDA, "%s: graph capture failed: %s", what, cudaGetErrorString(e));
    e = cudaGraphInstantiate(&slot->exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) { slot->exec = nullptr; return set_error(ECHR_ERR_CUDA, "%s: graph instantiate failed: %s", what, cudaGetErrorString(e)); }
    if (traced) b->trace_on = b->trace_on ? 2 : 0;  // 2: slots assigned, no more trace_next() calls
    slot->kernels = g_launches.load() - before;
    g_launches.store(before);
    memcpy(slot->key, key, sizeof(slot->key));
  }
  slot->stamp = ++b->graph_clock;
  if (traced && b->trace_on) {  // reset {min start, max end} before the replay
    ECHR_CUDA(cudaMemsetAsync(b->trace, 0xff, sizeof(unsigned long long) * TRACE_SLOTS * 2, st));
    ECHR_CUDA(cudaMemsetAsync(b->trace_detail, 0, sizeof(unsigned long long) * TRACE_SLOTS * 4, st));
  }
  ECHR_CUDA(cudaGraphLaunch(slot->exec, st));
  count_launch((int)slot->kernels);
  return ECHR_OK;
}

// =============================================================================================
// Encode: context prologue + relation encoder
// =============================================================================================
// clip planes + hoisted ctx2att projection over every clip of the batch (OldModel_NEW.py:381,
// computed once instead of once per decode step); the table holds e^{2P} (see kernels_step.cu)
static int encode_clips(echr_batch_t* b, cudaStream_t st) {
  const echr_model* m = b->model;
  const int at = m->cfg.a_terms, wt = m->cfg.w_terms;
  ECHR_CUDA(launch_split(b->c3d, b->n_clips, ECHR_VIDEO_DIM, ECHR_VIDEO_DIM, b->clip_pl.plane(0),
                         at > 1 ? b->clip_pl.plane(1) : nullptr, b->clip_pl.ld, CLIP_PAD, at == 1, st));
  GemmSpec s;
  s.A = &b->clip_pl; s.M = b->n_clips; s.W = &m->w_ctx2att; s.N = HID; s.K = CLIP_PAD;
  s.bias = m->b_ctx2att;
  // two-plane precisions: the fp32 table e^{2P}; single-plane: P, then the row-scaled fp16 table (attention_f16.cuh)
  s.out_f32 = b->proj_f32; s.ld_f32 = HID;
  if (at != 1) s.act = ACT_EXP2X;
  ECHR_TRY(run_gemm(&s, 1, at, wt, 128, EPI_STORE, 1, st, "ctx2att"));
  if (at == 1) ECHR_CUDA(launch_attention_table(b->proj_f32, b->n_clips, b->proj_h, 2 * HID, b->proj_scale, st));
  return ECHR_OK;
}

// event relation encoder from the event feature planes (MA:35-177): embedding -> query | key |
// value*W_out, pair position embedding -> fc1 + tanh -> fc2, multi-head relation attention
// `side` (optional, with the two events): the pair-position chain (embedding -> fc1 -> fc2) runs next to the
// event chain (embedding -> query | key | value); they meet at the relation attention.
static int encode_relation(echr_batch_t* b, cudaStream_t st, cudaStream_t side = nullptr, cudaEvent_t fork_ev = nullptr,
                           cudaEvent_t join_ev = nullptr) {
  const echr_model* m = b->model;
  const int at = m->cfg.a_terms, wt = m->cfg.w_terms;
  cudaStream_t ps = side != nullptr ? side : st;  // stream of the pair-position chain
  if (side != nullptr) {
    ECHR_CUDA(cudaEventRecord(fork_ev, st));
    ECHR_CUDA(cudaStreamWaitEvent(side, fork_ev, 0));
  }
  {
    GemmSpec s;
    s.A = &b->feat_pl; s.M = b->n_events; s.W = &m->w_event_emb; s.N = HID; s.K = FEAT_PAD;
    s.bias = m->b_event_emb;
    s.out_hi = b->x_pl.plane(0); s.out_lo = at > 1 ? b->x_pl.plane(1) : nullptr; s.ld_bf = HID;
    ECHR_TRY(run_gemm(&s, 1, at, wt, 128, EPI_STORE, 1, st, "event_emb"));
  }
  {
    GemmSpec s;
    s.A = &b->x_pl; s.M = b->n_events; s.W = &m->w_qky; s.N = 3 * HID; s.K = HID;
    s.bias = m->b_qky; s.out_f32 = b->qky; s.ld_f32 = 3 * HID;
    ECHR_TRY(run_gemm(&s, 1, at, wt, 128, EPI_STORE, 1, st, "query_key_value"));
  }
  // pair chain: position embedding -> fc1 + tanh -> fc2 in one kernel, nothing of size [pairs,512] touches HBM
  ECHR_TRY(launch_pair_mlp(b, ps));
  if (side != nullptr) {
    ECHR_CUDA(cudaEventRecord(join_ev, side));
    ECHR_CUDA(cudaStreamWaitEvent(st, join_ev, 0));
  }
  ECHR_CUDA(launch_relation_attn(b, st));
  return ECHR_OK;
}

// The encode pipeline as three concurrent chains: clips (plane split -> ctx2att), events, event pairs.
static int enqueue_encode(echr_batch_t* b, cudaStream_t st, cudaStream_t s2, cudaStream_t s3, cudaEvent_t* ev) {
  ECHR_CUDA(cudaEventRecord(ev[0], st));
  ECHR_CUDA(cudaStreamWaitEvent(s2, ev[0], 0));
  ECHR_TRY(encode_clips(b, s2));
  ECHR_CUDA(cudaEventRecord(ev[1], s2));
  if (b->n_events > 0) {
    ECHR_CUDA(launch_event_feats(b, st));
    ECHR_TRY(encode_relation(b, st, s3, ev[2], ev[3]));
  }
  ECHR_CUDA(cudaStreamWaitEvent(st, ev[1], 0));
  return ECHR_OK;
}

extern "C" int echr_encode(echr_batch_t* b, void* stream) {
  ECHR_REQUIRE(b, "echr_encode: null batch");
  ECHR_REQUIRE(b->n_videos > 0, "echr_encode: no inputs set");
  cudaStream_t st = (cudaStream_t)stream;
  static const bool graphed = !(knob("ECHR_ENCODE_GRAPH") && knob("ECHR_ENCODE_GRAPH")[0] == '0');
  const long long key[6] = {2, b->n_videos, b->n_events, b->n_clips, b->n_pairs, b->max_n};
  const bool use_graph = graphed && worth_a_graph(b, key);
  if (use_graph) {
    ECHR_TRY(run_graph(b, key, st, "encode", false, [&](cudaStream_t cs) {
      return enqueue_encode(b, cs, b->cap_stream2, b->cap_stream3, b->cap_events);
    }));
  } else {
    // eager: the same three chains on the batch's side streams
    ECHR_TRY(ensure_capture_streams(b));
    ECHR_TRY(enqueue_encode(b, st, b->cap_stream2, b->cap_stream3, b->cap_events));
  }
  b->encoded = true;
  return ECHR_OK;
}

// MA_Attention8.forward(feats, soi_select_list) on its own (models/MA_attention_8_NEW.py:35-49):
// feats fp32 device [n,1012] = (pooled clip feature | proposal state) per event, host_soi int32 [n,2]
// half-open spans of ONE video -> out fp32 device [n,512].
extern "C" int echr_relation_encode(echr_batch_t* b, const float* feats, const int32_t* host_soi, int n,
                                    float* out, void* stream) {
  ECHR_REQUIRE(b && feats && host_soi && out, "echr_relation_encode: null argument");
  ECHR_REQUIRE(n >= 1 && n <= b->cap.max_events && (long long)n * n <= b->cap.max_pairs,
               "echr_relation_encode: %d events exceed the batch capacity", n);
  cudaStream_t st = (cudaStream_t)stream;
  std::vector<int32_t> ev_video(n, 0), ev_s(n), ev_e(n), ev_ind(n, 0);
  int32_t T = 1;
  for (int i = 0; i < n; ++i) {
    ev_s[i] = host_soi[2 * i];
    ev_e[i] = host_soi[2 * i + 1];
    if (ev_e[i] > T) T = ev_e[i];
  }
  const int32_t offs[2] = {0, T};
  echr_batch_desc_t d = {1, n, offs, ev_video.data(), ev_s.data(), ev_e.data(), ev_ind.data()};
  ECHR_TRY(stage_desc(b, &d, false, st));
  ECHR_CUDA(launch_split(feats, n, ECHR_VIDEO_DIM + HID, ECHR_VIDEO_DIM + HID, b->feat_pl.plane(0),
                         b->feat_pl.planes > 1 ? b->feat_pl.plane(1) : nullptr, FEAT_PAD, FEAT_PAD, b->feat_pl.planes == 1, st));
  ECHR_TRY(encode_relation(b, st));
  ECHR_CUDA(cudaMemcpyAsync(out, b->event_ctx, sizeof(float) * (size_t)n * HID, cudaMemcpyDeviceToDevice, st));
  return ECHR_OK;
}

// lm_model-level entry (models/OldModel_NEW.py:98,139: forward / sample take ready-made contexts):
// video fp32 [100], event fp32 [n,512], clip fp32 [n,lmax,500] zero padded, host_clip_len int32 [n] =
// number of valid clips per event (the reference's clip_mask is a prefix mask).  After this call the
// decode entry points run on these contexts.
extern "C" int echr_batch_set_context(echr_batch_t* b, const float* video, const float* event, const float* clip,
                                      const int32_t* host_clip_len, int n, int lmax, void* stream) {
  ECHR_REQUIRE(b && video && event && clip && host_clip_len, "echr_batch_set_context: null argument");
  ECHR_REQUIRE(n >= 1 && n <= b->cap.max_events && lmax >= 1 && (long long)n * lmax <= b->cap.max_clips,
               "echr_batch_set_context: %d events x %d clips exceed the batch capacity", n, lmax);
  cudaStream_t st = (cudaStream_t)stream;
  // one "video" whose clip table is the padded clip tensor itself: event i owns rows [i*lmax, i*lmax+len)
  std::vector<int32_t> ev_video(n, 0), ev_s(n), ev_e(n), ev_ind(n, 0);
  for (int i = 0; i < n; ++i) {
    ECHR_REQUIRE(host_clip_len[i] >= 1 && host_clip_len[i] <= lmax, "echr_batch_set_context: clip_len[%d]=%d out of range", i, host_clip_len[i]);
    ev_s[i] = i * lmax;
    ev_e[i] = i * lmax + host_clip_len[i];
  }
  const int32_t offs[2] = {0, n * lmax};
  echr_batch_desc_t d = {1, n, offs, ev_video.data(), ev_s.data(), ev_e.data(), ev_ind.data()};
  ECHR_TRY(stage_desc(b, &d, true, st));
  ECHR_CUDA(cudaMemcpyAsync(b->c3d, clip, sizeof(float) * (size_t)n * lmax * ECHR_VIDEO_DIM, cudaMemcpyDeviceToDevice, st));
  ECHR_CUDA(cudaMemcpyAsync(b->lda, video, sizeof(float) * ECHR_LDA_DIM, cudaMemcpyDeviceToDevice, st));
  ECHR_CUDA(cudaMemcpyAsync(b->event_ctx, event, sizeof(float) * (size_t)n * HID, cudaMemcpyDeviceToDevice, st));
  ECHR_TRY(encode_clips(b, st));
  b->encoded = true;
  return ECHR_OK;
}

extern "C" int echr_get_event_context(const echr_batch_t* b, float* out, void* stream) {
  ECHR_REQUIRE(b && out, "echr_get_event_context: null argument");
  ECHR_REQUIRE(b->encoded, "echr_get_event_context: call echr_encode first");
  ECHR_CUDA(cudaMemcpyAsync(out, b->event_ctx, sizeof(float) * (size_t)b->n_events * HID,
                            cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return ECHR_OK;
}

// =============================================================================================
// Decode
// =============================================================================================
// k-block tables over the activation matrix, in the weights' column order [x | h | ctx]
// (the x block is not part of the step GEMM: it is the token gate table, gathered in the epilogue)
static const int KB_L0[8] = {26, 27, 28, 29, 30, 31, 32, 33};                                   // h0
static const int KB_L1[16] = {34, 35, 36, 37, 38, 39, 40, 41, 16, 17, 18, 19, 20, 21, 22, 23};  // h1, att
static const int KB_L2[8] = {42, 43, 44, 45, 46, 47, 48, 49};                                   // h2
// with the constant-context blocks kept in the step GEMM (weights' column order [x | h | ctx])
static const int KB_L0C[16] = {26, 27, 28, 29, 30, 31, 32, 33, 8, 9, 10, 11, 12, 13, 14, 15};   // h0, event
static const int KB_L2C[10] = {42, 43, 44, 45, 46, 47, 48, 49, 24, 25};                         // h2, video

enum { MODE_GREEDY = 0, MODE_BEAM = 1, MODE_TF = 2 };

static int enqueue_init(echr_batch* b, int beam, cudaStream_t st) {
  const long long rows = (long long)b->n_events * beam;
  for (int i = 0; i < 2; ++i)
    for (int p = 0; p < b->act[i].planes; ++p)
      ECHR_CUDA(cudaMemsetAsync(b->act[i].plane(p), 0, sizeof(bf16) * (size_t)rows * ACT_LD, st));
  ECHR_CUDA(cudaMemsetAsync(b->cstate[0], 0, sizeof(float) * (size_t)3 * rows * HID, st));
  ECHR_CUDA(launch_decode_init(b, beam, st));
  // Hoist the constant-context terms of streams 0 and 2 out of the time loop: per-row gate bias
  //   row_bias0 = W_ih0[:, event] . event_ctx + (b_ih0 + b_hh0),  row_bias2 = W_ih2[:, video] . lda + (b_ih2 + b_hh2)
  // (ThreeStream_Core feeds the same event / video context at every step, OldModel_NEW.py:806-817).
  if (!hoist_context((int)rows)) return ECHR_OK;
  const echr_model* m = b->model;
  const int at = m->cfg.a_terms, wt = m->cfg.w_terms;
  GemmSpec s[2];
  PlaneMat wv[2];
  const int layer[2] = {0, 2};
  const int col0[2] = {ACT_EV, ACT_VID};
  const int K[2] = {HID, VID_PAD};
  for (int i = 0; i < 2; ++i) {
    wv[i] = m->w_lstm[layer[i]];
    wv[i].p += 2 * HID;  // the context block of [x | h | ctx]; plane stride and ld are unchanged
    s[i].A = &b->act[0]; s[i].a_col0 = col0[i]; s[i].M = (int)rows; s[i].W = &wv[i]; s[i].N = GATES; s[i].K = K[i];
    s[i].bias = nullptr;  // the fused bias is part of the token gate tables
    s[i].out_f32 = b->row_bias[i]; s[i].ld_f32 = GATES;
  }
  ECHR_TRY(run_gemm(s, 2, at, wt, 128, EPI_STORE, 1, st, "hoisted_context"));  // both streams in one launch
  return ECHR_OK;
}

// One decode step t: attention query + attention (stream 1 uses the PREVIOUS h1), the three gate
// GEMMs with the fused LSTM cell update, the vocabulary projection with fused partial
// log-softmax / top-k.  The caller appends the mode's finalize kernel.
// PART_NEXT_Q (with PART_LOGIT, persistent kernel only): the attention query GEMM of step t + 1 — it needs this
// step's h1 and nothing of the token — is a second problem of the logit launch of step t.
enum { PART_H2ATT = 1, PART_ATT = 2, PART_GATES = 4, PART_LOGIT = 8, PART_ALL = 15, PART_NEXT_Q = 16 };

// Attention of decode step t over the 16-bit tables (single-plane precision): reads the query the h2att GEMM of
// step t wrote, writes the ATT block of activation buffer t & 1.
static AttF16Params attention_params(const echr_batch* b, int rows, int t) {
  AttF16Params ap;
  ap.q_exp = b->q_att;
  ap.tab = b->proj_h;
  ap.proj_scale = b->proj_scale;
  ap.alpha = b->model->alpha;
  ap.row_start = b->row_start;
  ap.row_len = b->row_len;
  ap.rows = rows;
  ap.out = b->act[t & 1].plane(0) + ACT_ATT;
  ap.ld_out = ACT_LD;
  ap.out_f32 = nullptr;
  return ap;
}

// use_keys (greedy decode): the logit epilogue keeps a packed running arg-max per row (argmax_key[t & 1]) and the
// gate GEMM of step t + 1 takes its input token from it, so the token chain does not pass through finalize.
// c_src (beam search): the gate epilogue takes row r's previous cell state from row c_src[r] (the parent slot).
static int enqueue_step(echr_batch* b, int rows, int t, int topk, float* logits_out, long long ld_logits,
                        cudaStream_t st, int parts = PART_ALL, bool use_keys = false, const int32_t* c_src = nullptr) {
  const echr_model* m = b->model;
  const int at = m->cfg.a_terms, wt = m->cfg.w_terms;
  const int p = t & 1, n = p ^ 1;
  PlaneMat& cur = b->act[p];
  PlaneMat& nxt = b->act[n];
  const bool att16 = at == 1;  // single-plane precision: 16-bit tables, attention_f16.cuh
  if (parts & PART_H2ATT) {
    GemmSpec s;
    s.A = &cur; s.a_col0 = ACT_H + HID; s.M = rows; s.W = &m->w_h2att; s.N = HID; s.K = HID;
    s.bias = m->b_h2att; s.act = ACT_EXP2X; s.out_f32 = b->q_att; s.ld_f32 = HID;  // e^{2q}
    PdlScope pdl(PDL_H2ATT);
    // latency regime: narrow tiles spread the [rows,512] x [512,512] product over 4x the SMs (a CTA's
    // main loop is bound by its own L2->SM operand stream, dominated by the 2-plane activation tile)
    // (lanes sharing the GPU: one wide tile per 128 rows costs a third of the SM time of four narrow ones)
    // (throughput regime: 128 x 64 tiles, one per CTA: a [rows,512] x [512,512] product has only rows/64 tiles of
    // the persistent kernel's 128 x 256 shape, 32 CTAs at 2048 rows)
    // (beam search over thousands of events — from half a 128 x 256 tile per SM on — the persistent kernel; measured
    // us per launch, 128 x 64 tiles / persistent: 1166 rows 6.9 / 11.9, 2045: 6.8 / 11.6, 4072: 12.6 / 11.9, 9600: 29.3 / 19.4)
    const int bn = rows < gemm2_min_rows() ? (b->concurrency > 1 ? 128 : 32) : rows >= 32 * num_sms() ? 128 : 64;
    ECHR_TRY(run_gemm(&s, 1, at, wt, bn, EPI_STORE, 1, st, "h2att", b, 0));
  }
  if (parts & PART_ATT) {
    PdlScope pdl(PDL_ATT);
    if (att16) {
      const AttF16Params ap = attention_params(b, rows, t);
      ECHR_CUDA(launch_attention_f16(ap, b->mean_len, st, trace_ptr(b), trace_next(b, 1)));
    } else {
      ECHR_CUDA(launch_attention(b->q_att, (const void*)b->proj_f32, (const void*)b->c3d, ECHR_VIDEO_DIM, 0, m->alpha,
                                 b->row_start, b->row_len, rows, nullptr, cur.plane(0) + ACT_ATT,
                                 at > 1 ? cur.plane(1) + ACT_ATT : nullptr, ACT_LD, st, len_bucket(b->max_len), trace_ptr(b),
                                 trace_next(b, 1)));
    }
  }
  if (parts & PART_GATES) {
    GemmSpec s[3];
    PlaneMat wv[3];
    // latency regime: stream 1 has twice the k-blocks of the others; give its tiles to CTA pairs (and the
    // other streams' too while everything still fits one wave)
    int ks[3] = {1, 1, 1};
    if ((ksplit_mask() & 1) != 0 && b->concurrency <= 1 && rows < gemm2_min_rows()) {
      const int mt = m_tiles_of(rows), nt = GATES / 128;
      if (mt * nt * 6 <= num_sms()) ks[0] = ks[1] = ks[2] = 2;
      else if (mt * nt * 4 <= num_sms()) ks[1] = 2;
    }
    const bool hoist = hoist_context(rows);
    const int* kbs[3] = {hoist ? KB_L0 : KB_L0C, KB_L1, hoist ? KB_L2 : KB_L2C};
    const int K[3] = {(hoist ? 8 : 16) * 64, 16 * 64, (hoist ? 8 : 10) * 64};
    for (int l = 0; l < 3; ++l) {
      wv[l] = m->w_lstm[l];
      wv[l].p += HID;  // skip the x block of [x | h | ctx]; plane stride and ld are unchanged
      s[l].A = &cur; s[l].a_kb = kbs[l]; s[l].M = rows; s[l].W = &wv[l]; s[l].N = GATES; s[l].K = K[l];
      s[l].bias = m->b_lstm[l];
      if (l != 1 && hoist) { s[l].row_bias = b->row_bias[l / 2]; s[l].ld_row_bias = GATES; }
      s[l].row_tok = b->row_tok; s[l].tok_table = m->tok_gates[l]; s[l].ld_tok_table = GATES;
      if (use_keys) s[l].row_key = b->argmax_key[(t + 1) & 1];  // written by the logits of step t - 1 (BOS at t = 0)
      s[l].ksplit = ks[l];
      s[l].c_in = b->cstate[p] + (long long)l * rows * HID;
      s[l].c_src = c_src;
      s[l].c_out = b->cstate[n] + (long long)l * rows * HID;
      s[l].out_hi = nxt.plane(0) + ACT_H + l * HID;
      s[l].out_lo = at > 1 ? nxt.plane(1) + ACT_H + l * HID : nullptr;
      s[l].ld_bf = ACT_LD;
    }
    PdlScope pdl(PDL_GATES);
    ECHR_TRY(run_gemm(s, 3, at, wt, 128, EPI_LSTM, 1, st, "lstm_gates", b, 2));
  }
  if (parts & PART_LOGIT) {
    GemmSpec sv[2];
    GemmSpec& s = sv[0];
    s.A = &nxt; s.a_col0 = ACT_H; s.M = rows; s.W = &m->w_logit; s.N = m->V1; s.K = 3 * HID;
    s.bias = m->b_logit; s.partial = b->partial; s.logits_out = logits_out; s.ld_logits = ld_logits;
    const bool sp = logit_split(m, rows);
    if (sp) s.ksplit = 2;
    if (use_keys) s.argmax_key = b->argmax_key[t & 1];
    int groups = 1;
    if (parts & PART_NEXT_Q) {
      // the short tiles go last: they fill the tail of the persistent kernel's last round of vocabulary tiles
      GemmSpec& q = sv[1];
      q.A = &nxt; q.a_col0 = ACT_H + HID; q.M = rows; q.W = &m->w_h2att; q.N = HID; q.K = HID;
      q.bias = m->b_h2att; q.act = ACT_EXP2X; q.out_f32 = b->q_att; q.ld_f32 = HID; q.epi_store = true;
      groups = 2;
    }
    PdlScope pdl(PDL_LOGIT);
    ECHR_TRY(run_gemm(sv, groups, at, wt, sp ? 256 : 128, EPI_LOGITS, topk, st, "logit", b, 3));
  }
  return ECHR_OK;
}

static int enqueue_decode(echr_batch* b, int beam, int mode, cudaStream_t st, cudaStream_t side = nullptr) {
  const int rows = b->n_events * beam;
  const int S = b->model->cfg.seq_length;
  const int topk = mode == MODE_BEAM ? 4 : 1;
  // programmatic dependent launches hide launch latency and prologues, which only matter while a step is
  // tens of microseconds (measured: +9 % events/s at 251 rows, -1 % at 4072)
  struct Regime {
    bool prev;
    explicit Regime(bool on) : prev(pdl_regime()) { pdl_regime() = on; }
    ~Regime() { pdl_regime() = prev; }
  } regime(rows < gemm2_min_rows());
  ECHR_TRY(enqueue_init(b, beam, st));
  if (mode == MODE_BEAM) {
    ECHR_CUDA(cudaMemsetAsync(b->tokens_dev, 0, sizeof(int32_t) * (size_t)b->n_events * S, st));
    ECHR_CUDA(cudaMemsetAsync(b->logp_dev, 0, sizeof(float) * (size_t)b->n_events * S, st));
  }
  if (mode == MODE_GREEDY && side != nullptr) {
    // Greedy: the next step's attention (h2att GEMM + fused attention) depends only on this step's h1, not
    // on the token, so it runs next to the vocabulary projection.  The recurrent chain
    //   gates(t) -> h2att(t+1) -> attention(t+1) -> gates(t+1)
    // is the critical one and stays on one stream, so each of its kernels is a programmatic dependent of the
    // previous one (launch latency and prologue hidden behind the predecessor's epilogue); the token chain
    //   gates(t) -> logits(t) -> finalize(t) -> gates(t+1)
    // forks off and joins by events.  The branches touch disjoint columns of the activation matrix.
    // (throughput regime: persistent GEMMs; measured better with the attention chain on the forked stream)
    const bool crit_main = rows < gemm2_min_rows();
    if (!crit_main && fuse_next_query() && use_gemm2(rows, b->model->V1, 128)) {  // (a vocabulary below one tile: not persistent)
      // Throughput regime: every kernel fills the GPU, so nothing is gained by running two of them side by side;
      // what counts is the number of launches and their tails.  The attention query GEMM of step t + 1 (a
      // [rows,512] x [512,512] product: 12.7 us as its own launch at 4072 rows, mostly ramp and tail) becomes
      // the last tiles of the logit launch of step t, and finalize(t) — off the token chain thanks to the
      // arg-max keys — runs on the forked stream beside attention(t + 1) and gates(t + 1):
      //   st:   gates(t) -> [logits(t) | query(t+1)] -> attention(t+1) -> gates(t+1) -> ...
      //   side:                                      -> finalize(t) ----------------------^ (before logits(t+1):
      //         it reads the partial records and clears the keys that launch accumulates into)
      ECHR_TRY(enqueue_step(b, rows, 0, topk, nullptr, 0, st, PART_H2ATT | PART_ATT));
      for (int t = 0; t < S; ++t) {
        const bool last = t + 1 == S;
        ECHR_TRY(enqueue_step(b, rows, t, topk, nullptr, 0, st, PART_GATES, true));
        if (t > 0) ECHR_CUDA(cudaStreamWaitEvent(st, b->cap_events[1], 0));  // finalize(t - 1) done
        ECHR_TRY(enqueue_step(b, rows, t, topk, nullptr, 0, st, PART_LOGIT | (last ? 0 : PART_NEXT_Q), true));
        ECHR_CUDA(cudaEventRecord(b->cap_events[0], st));
        ECHR_CUDA(cudaStreamWaitEvent(side, b->cap_events[0], 0));
        ECHR_CUDA(launch_finalize_greedy(b, t, b->tokens_dev, b->logp_dev, side, true));
        ECHR_CUDA(cudaEventRecord(b->cap_events[1], side));
        if (!last) ECHR_TRY(enqueue_step(b, rows, t + 1, topk, nullptr, 0, st, PART_ATT));
      }
      ECHR_CUDA(cudaStreamWaitEvent(st, b->cap_events[1], 0));
      return ECHR_OK;
    }
    // The gate GEMM of step t+1 reads its token from the arg-max keys the logit epilogue of step t maintains
    // with atomics (ECHR_ARGMAX_KEYS=0: from finalize's row_tok), so the join waits for the logits only;
    // finalize (log-sum-exp, outputs, `unfinished`) runs beside the next gate GEMM.
    const bool keys = true;
    ECHR_TRY(enqueue_step(b, rows, 0, topk, nullptr, 0, st, PART_H2ATT | PART_ATT));
    cudaStream_t tok_st = crit_main ? side : st;  // token chain
    cudaStream_t att_st = crit_main ? st : side;  // recurrent attention chain
    for (int t = 0; t < S; ++t) {
      ECHR_TRY(enqueue_step(b, rows, t, topk, nullptr, 0, st, PART_GATES, keys));
      ECHR_CUDA(cudaEventRecord(b->cap_events[0], st));
      ECHR_CUDA(cudaStreamWaitEvent(side, b->cap_events[0], 0));
      if (!crit_main && t + 1 < S) {
        ECHR_TRY(enqueue_step(b, rows, t + 1, topk, nullptr, 0, att_st, PART_H2ATT | PART_ATT));
        ECHR_CUDA(cudaEventRecord(b->cap_events[1], att_st));
      }
      ECHR_TRY(enqueue_step(b, rows, t, topk, nullptr, 0, tok_st, PART_LOGIT, keys));
      const bool last = t + 1 == S;
      if (keys && !last) ECHR_CUDA(cudaEventRecord(b->cap_events[2], tok_st));  // logits(t) done
      {
        PdlScope pdl(PDL_FINALIZE);
        ECHR_CUDA(launch_finalize_greedy(b, t, b->tokens_dev, b->logp_dev, tok_st, keys));
      }
      if (crit_main) {
        if (!keys || last) ECHR_CUDA(cudaEventRecord(b->cap_events[2], tok_st));  // finalize(t) done
        if (!last) ECHR_TRY(enqueue_step(b, rows, t + 1, topk, nullptr, 0, att_st, PART_H2ATT | PART_ATT));
        ECHR_CUDA(cudaStreamWaitEvent(st, b->cap_events[2], 0));
      } else {
        // token chain on st itself: with keys the next gates would still follow finalize in stream order, so
        // this arrangement gains nothing from them; join the attention chain
        if (!last) ECHR_CUDA(cudaStreamWaitEvent(st, b->cap_events[1], 0));
      }
    }
    return ECHR_OK;
  }
  for (int t = 0; t < S; ++t) {
    // beam: finalize_beam(t - 1) reordered h (a TMA operand has to be in place) but left the cell state where it
    // was and wrote each row's parent row instead
    ECHR_TRY(enqueue_step(b, rows, t, topk, nullptr, 0, st, PART_ALL, false, mode == MODE_BEAM && t > 0 ? b->row_parent : nullptr));
    if (mode == MODE_BEAM)
      ECHR_CUDA(launch_finalize_beam(b, beam, t, b->tokens_dev, b->logp_dev, b->score_dev, st));
    else
      ECHR_CUDA(launch_finalize_greedy(b, t, b->tokens_dev, b->logp_dev, st));
  }
  return ECHR_OK;
}

// The whole decode (init + seq_length steps) is one CUDA graph per (rows, beam mode, window bucket): one
// launch from the host per batch instead of ~150.
static int run_decode(echr_batch* b, int beam, int mode, cudaStream_t st) {
  ECHR_REQUIRE(b->encoded, "decode: call echr_encode first");
  ECHR_REQUIRE((long long)b->n_events * beam <= b->cap.max_rows, "decode: %d events x beam %d exceed max_rows %lld", b->n_events, beam, (long long)b->cap.max_rows);
  if (b->n_events == 0) return ECHR_OK;
  const int rows = b->n_events * beam;
  const long long key[6] = {1, rows, mode == MODE_BEAM ? beam : 0, len_bucket(b->max_len), b->concurrency > 1 ? 1 : 0, 0};
  const char* tr = knob("ECHR_TRACE");
  if (!(tr != nullptr && tr[0] == '1') && !worth_a_graph(b, key)) {
    // first sight of this shape: the same fork/join schedule, enqueued eagerly on the caller's stream
    ECHR_TRY(ensure_capture_streams(b));
    return enqueue_decode(b, beam, mode, st, b->cap_stream2);
  }
  return run_graph(b, key, st, "decode", true,
                   [&](cudaStream_t cs) { return enqueue_decode(b, beam, mode, cs, b->cap_stream2); });
}

static int copy_results(echr_batch* b, int32_t* tokens, float* logp, float* score, cudaMemcpyKind kind,
                        cudaStream_t st) {
  const size_t n = (size_t)b->n_events * b->model->cfg.seq_length;
  if (n == 0) return ECHR_OK;
  if (tokens) ECHR_CUDA(cudaMemcpyAsync(tokens, b->tokens_dev, sizeof(int32_t) * n, kind, st));
  if (logp) ECHR_CUDA(cudaMemcpyAsync(logp, b->logp_dev, sizeof(float) * n, kind, st));
  if (score) ECHR_CUDA(cudaMemcpyAsync(score, b->score_dev, sizeof(float) * (size_t)b->n_events, kind, st));
  return ECHR_OK;
}

extern "C" int echr_decode_greedy(echr_batch_t* b, int32_t* tokens, float* logp, void* stream) {
  ECHR_REQUIRE(b, "echr_decode_greedy: null batch");
  cudaStream_t st = (cudaStream_t)stream;
  ECHR_TRY(run_decode(b, 1, MODE_GREEDY, st));
  return copy_results(b, tokens, logp, nullptr, cudaMemcpyDeviceToDevice, st);
}

extern "C" int echr_decode_beam(echr_batch_t* b, int beam, int32_t* tokens, float* logp, float* score,
                                void* stream) {
  ECHR_REQUIRE(b, "echr_decode_beam: null batch");
  ECHR_REQUIRE(beam >= 1 && beam <= 4, "echr_decode_beam: beam must be in 1..4");
  cudaStream_t st = (cudaStream_t)stream;
  ECHR_TRY(run_decode(b, beam, MODE_BEAM, st));
  return copy_results(b, tokens, logp, score, cudaMemcpyDeviceToDevice, st);
}

// The decisions of the last echr_decode_beam / beam caption call: for every step t and decode row r = event * beam +
// slot, the parent slot (within the event) and the token the new slot r was built from.  parent / tok: int32 device
// [seq_length, rows].  This is what the parity tests replay with the oracle's arithmetic.
extern "C" int echr_get_beam_trace(const echr_batch_t* b, int beam, int32_t* parent, int32_t* tok, void* stream) {
  ECHR_REQUIRE(b && parent && tok, "echr_get_beam_trace: null argument");
  ECHR_REQUIRE(beam >= 1 && (long long)b->n_events * beam <= b->cap.max_rows, "echr_get_beam_trace: bad beam");
  const int S = b->model->cfg.seq_length;
  const size_t rows = (size_t)b->n_events * beam;
  if (rows == 0) return ECHR_OK;
  cudaStream_t st = (cudaStream_t)stream;
  ECHR_CUDA(cudaMemcpy2DAsync(parent, rows * sizeof(int32_t), b->beam_trace, (size_t)b->cap.max_rows * sizeof(int32_t),
                              rows * sizeof(int32_t), S, cudaMemcpyDeviceToDevice, st));
  ECHR_CUDA(cudaMemcpy2DAsync(tok, rows * sizeof(int32_t), b->beam_trace + (size_t)S * b->cap.max_rows,
                              (size_t)b->cap.max_rows * sizeof(int32_t), rows * sizeof(int32_t), S, cudaMemcpyDeviceToDevice, st));
  return ECHR_OK;
}

extern "C" int echr_teacher_forced(echr_batch_t* b, const int32_t* seq, int seq_cols, float* logprobs,
                                   void* stream) {
  ECHR_REQUIRE(b && seq && logprobs, "echr_teacher_forced: null argument");
  ECHR_REQUIRE(b->encoded, "echr_teacher_forced: call echr_encode first");
  ECHR_REQUIRE(seq_cols >= 2 && seq_cols <= 65, "echr_teacher_forced: seq_cols must be in 2..65");
  cudaStream_t st = (cudaStream_t)stream;
  if (b->n_events == 0) return ECHR_OK;
  const int rows = b->n_events;
  const int V1 = b->model->V1;
  ECHR_TRY(enqueue_init(b, 1, st));
  ECHR_CUDA(launch_finalize_tf(b, -1, seq, seq_cols, logprobs, st));
  for (int t = 0; t < seq_cols - 1; ++t) {
    ECHR_TRY(enqueue_step(b, rows, t, 1, logprobs + (long long)t * V1, (long long)(seq_cols - 1) * V1, st));
    ECHR_CUDA(launch_finalize_tf(b, t, seq, seq_cols, logprobs, st));
  }
  return ECHR_OK;
}

// The teacher-forced pass as the loss needs it (eval_utils.py:139-152, train.py:298-300: cg_crit(cg_model(...,
// mode='train'), labels[:,1:], masks[:,1:])): instead of the [n, S-1, V+1] log-prob tensor (173 MB at 251 events)
// only target_logp[r, t] = log p(seq[r, t+1] | seq[r, :t+1]) is produced, [n_events, seq_cols-1]: exactly the
// entries LanguageModelCriterion gathers (misc/utils.py:62-75).
extern "C" int echr_teacher_forced_nll(echr_batch_t* b, const int32_t* seq, int seq_cols, float* target_logp,
                                       void* stream) {
  ECHR_REQUIRE(b && seq && target_logp, "echr_teacher_forced_nll: null argument");
  ECHR_REQUIRE(b->encoded, "echr_teacher_forced_nll: call echr_encode first");
  ECHR_REQUIRE(seq_cols >= 2 && seq_cols <= 65, "echr_teacher_forced_nll: seq_cols must be in 2..65");
  cudaStream_t st = (cudaStream_t)stream;
  if (b->n_events == 0) return ECHR_OK;
  const int rows = b->n_events;
  const int V1 = b->model->V1;
  if (b->nll_rows < rows) {
    ECHR_CUDA(cudaStreamSynchronize(st));
    cudaFree(b->nll_logits);
    b->nll_logits = nullptr;
    b->nll_rows = 0;
    const long long want = std::max<long long>(rows, b->cap.max_events);
    if (cudaMalloc(&b->nll_logits, sizeof(float) * (size_t)want * V1) != cudaSuccess)
      return set_error(ECHR_ERR_WORKSPACE, "echr_teacher_forced_nll: cudaMalloc of %lld logit rows failed", want);
    b->nll_rows = want;
  }
  ECHR_TRY(enqueue_init(b, 1, st));
  ECHR_CUDA(launch_finalize_nll(b, -1, seq, seq_cols, b->nll_logits, target_logp, st));
  for (int t = 0; t < seq_cols - 1; ++t) {
    ECHR_TRY(enqueue_step(b, rows, t, 1, b->nll_logits, V1, st));
    ECHR_CUDA(launch_finalize_nll(b, t, seq, seq_cols, b->nll_logits, target_logp, st));
  }
  return ECHR_OK;
}

// OldModel.get_logprobs_state (OldModel_NEW.py:133-137) with explicit state: it int32 [n], h_in / c_in
// fp32 [3,n,512] -> logprobs fp32 [n,V+1], h_out / c_out fp32 [3,n,512].  Contexts come from
// echr_encode or echr_batch_set_context.  (h passes through the bf16 hi+lo planes: 16 mantissa bits.)
extern "C" int echr_decode_step(echr_batch_t* b, const int32_t* it, const float* h_in, const float* c_in,
                                float* logprobs, float* h_out, float* c_out, void* stream) {
  ECHR_REQUIRE(b && it && h_in && c_in && logprobs && h_out && c_out, "echr_decode_step: null argument");
  ECHR_REQUIRE(b->encoded && b->n_events > 0, "echr_decode_step: set the contexts first");
  cudaStream_t st = (cudaStream_t)stream;
  const int rows = b->n_events;
  const int V1 = b->model->V1;
  ECHR_TRY(enqueue_init(b, 1, st));
  ECHR_CUDA(launch_load_state(b, rows, it, h_in, st));
  ECHR_CUDA(cudaMemcpyAsync(b->cstate[0], c_in, sizeof(float) * (size_t)3 * rows * HID, cudaMemcpyDeviceToDevice, st));
  ECHR_TRY(enqueue_step(b, rows, 0, 1, logprobs, V1, st));
  ECHR_CUDA(launch_finalize_tf(b, 0, it, 2, logprobs, st));  // seq_cols = 2: normalise only
  ECHR_CUDA(launch_export_state(b, rows, h_out, st));
  ECHR_CUDA(cudaMemcpyAsync(c_out, b->cstate[1], sizeof(float) * (size_t)3 * rows * HID, cudaMemcpyDeviceToDevice, st));
  return ECHR_OK;
}

extern "C" int echr_batch_set_concurrency(echr_batch_t* b, int lanes) {
  ECHR_REQUIRE(b && lanes >= 1, "echr_batch_set_concurrency: bad argument");
  b->concurrency = lanes;
  return ECHR_OK;
}

static int caption_host(echr_batch_t* b, const echr_batch_desc_t* desc, const float* host_c3d,
                        const float* host_tap, const float* host_lda, int beam, int32_t* host_tokens,
                        float* host_logp, float* host_score, cudaStream_t st, bool sync) {
  ECHR_REQUIRE(b && host_tokens && host_logp, "echr_caption_host: null argument");
  ECHR_REQUIRE(beam <= 4, "echr_caption_host: beam must be in 1..4");
  ECHR_TRY(echr_batch_set_inputs(b, desc, host_c3d, host_tap, host_lda, 1, st));
  ECHR_TRY(echr_encode(b, st));
  ECHR_TRY(run_decode(b, beam > 1 ? beam : 1, beam > 1 ? MODE_BEAM : MODE_GREEDY, st));
  ECHR_TRY(copy_results(b, host_tokens, host_logp, beam > 1 ? host_score : nullptr, cudaMemcpyDeviceToHost, st));
  if (sync) ECHR_CUDA(cudaStreamSynchronize(st));
  return ECHR_OK;
}

extern "C" int echr_caption_host(echr_batch_t* b, const echr_batch_desc_t* desc, const float* host_c3d,
                                 const float* host_tap, const float* host_lda, int beam,
                                 int32_t* host_tokens, float* host_logp, float* host_score, void* stream) {
  return caption_host(b, desc, host_c3d, host_tap, host_lda, beam, host_tokens, host_logp, host_score,
                      (cudaStream_t)stream, true);
}

extern "C" int echr_caption_host_async(echr_batch_t* b, const echr_batch_desc_t* desc, const float* host_c3d,
                                       const float* host_tap, const float* host_lda, int beam,
                                       int32_t* host_tokens, float* host_logp, float* host_score, void* stream) {
  return caption_host(b, desc, host_c3d, host_tap, host_lda, beam, host_tokens, host_logp, host_score,
                      (cudaStream_t)stream, false);
}

extern "C" int echr_caption_device(echr_batch_t* b, const echr_batch_desc_t* desc, const float* c3d, const float* tap,
                                   const float* lda, int beam, int32_t* tokens, float* logp, float* score,
                                   void* stream) {
  ECHR_REQUIRE(b && tokens && logp, "echr_caption_device: null argument");
  ECHR_REQUIRE(beam >= 1 && beam <= 4, "echr_caption_device: beam must be in 1..4");
  cudaStream_t st = (cudaStream_t)stream;
  ECHR_TRY(echr_batch_set_inputs(b, desc, c3d, tap, lda, 0, st));
  ECHR_TRY(echr_encode(b, st));
  ECHR_TRY(run_decode(b, beam, beam > 1 ? MODE_BEAM : MODE_GREEDY, st));
  return copy_results(b, tokens, logp, beam > 1 ? score : nullptr, cudaMemcpyDeviceToDevice, st);
}

// Average duration (ms) of each kernel of one decode step, measured with CUDA events on `stream`
// around `iters` back-to-back launches of that kernel (state left by the previous decode).
extern "C" int echr_time_step_kernels(echr_batch_t* b, int beam, int iters, float* host_ms, void* stream) {
  ECHR_REQUIRE(b && host_ms && iters >= 1, "echr_time_step_kernels: bad argument");
  ECHR_REQUIRE(b->encoded && b->n_events > 0, "echr_time_step_kernels: decode a batch first");
  ECHR_REQUIRE(beam >= 1 && beam <= 4 && (long long)b->n_events * beam <= b->cap.max_rows, "echr_time_step_kernels: bad beam");
  cudaStream_t st = (cudaStream_t)stream;
  const int rows = b->n_events * beam;
  const long long before = g_launches.load();
  cudaEvent_t e0, e1;
  ECHR_CUDA(cudaEventCreate(&e0));
  ECHR_CUDA(cudaEventCreate(&e1));
  const int parts[4] = {PART_H2ATT, PART_ATT, PART_GATES, PART_LOGIT};
  int rc = enqueue_init(b, beam, st);  // fresh state (a finished beam decode leaves every slot dead)
  if (rc == ECHR_OK) rc = enqueue_step(b, rows, 0, beam > 1 ? 4 : 1, nullptr, 0, st);
  // ... and the bookkeeping of step 0, so that the kernels below see what step 1 of a decode sees: every row its
  // own input token (with the start token in all rows the gate GEMM's table gather would be one hot 8 KB row)
  if (rc == ECHR_OK) {
    const cudaError_t fe = beam > 1 ? launch_finalize_beam(b, beam, 0, b->tokens_dev, b->logp_dev, b->score_dev, st)
                                    : launch_finalize_greedy(b, 0, b->tokens_dev, b->logp_dev, st);
    if (fe != cudaSuccess) rc = set_error(ECHR_ERR_CUDA, "echr_time_step_kernels: finalize of step 0 failed");
  }
  // each kernel: `iters` launches captured into one graph (no host launch gaps between them, like inside the
  // decode graph), replayed once to warm up and once between the two events
  if (rc == ECHR_OK) rc = ensure_capture_streams(b);
  if (rc == ECHR_OK && cudaStreamSynchronize(st) != cudaSuccess) rc = set_error(ECHR_ERR_CUDA, "echr_time_step_kernels: sync failed");
  const bool saved_regime = pdl_regime();
  pdl_regime() = false;  // plain launches: a launch must not overlap the tail of its identical predecessor
  host_ms[5] = 0.f;
  const bool fused_q = beam == 1 && rows >= gemm2_min_rows() && fuse_next_query() && use_gemm2(rows, b->model->V1, 128);
  for (int k = 0; k < 6 && rc == ECHR_OK; ++k) {
    if (k == 5 && !fused_q) break;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    if (cudaStreamBeginCapture(b->cap_stream, cudaStreamCaptureModeRelaxed) != cudaSuccess) { rc = set_error(ECHR_ERR_CUDA, "echr_time_step_kernels: capture failed"); break; }
    for (int i = 0; i < iters && rc == ECHR_OK; ++i) {
      if (k < 4) rc = enqueue_step(b, rows, 1, beam > 1 ? 4 : 1, nullptr, 0, b->cap_stream, parts[k]);
      else if (k == 5) rc = enqueue_step(b, rows, 1, 1, nullptr, 0, b->cap_stream, PART_LOGIT | PART_NEXT_Q);
      else if (beam > 1) { if (launch_finalize_beam(b, beam, 1, b->tokens_dev, b->logp_dev, b->score_dev, b->cap_stream) != cudaSuccess) rc = set_error(ECHR_ERR_CUDA, "finalize_beam launch failed"); }
      else { if (launch_finalize_greedy(b, 1, b->tokens_dev, b->logp_dev, b->cap_stream) != cudaSuccess) rc = set_error(ECHR_ERR_CUDA, "finalize_greedy launch failed"); }
    }
    cudaError_t ce = cudaStreamEndCapture(b->cap_stream, &graph);
    if (rc == ECHR_OK && ce != cudaSuccess) rc = set_error(ECHR_ERR_CUDA, "echr_time_step_kernels: end capture: %s", cudaGetErrorString(ce));
    if (rc == ECHR_OK && cudaGraphInstantiate(&exec, graph, 0) != cudaSuccess) rc = set_error(ECHR_ERR_CUDA, "echr_time_step_kernels: instantiate failed");
    if (graph) cudaGraphDestroy(graph);
    for (int pass = 0; pass < 2 && rc == ECHR_OK; ++pass) {  // pass 0 warms up
      cudaEventRecord(e0, st);
      cudaGraphLaunch(exec, st);
      cudaEventRecord(e1, st);
      if (cudaEventSynchronize(e1) != cudaSuccess) rc = set_error(ECHR_ERR_CUDA, "echr_time_step_kernels: kernel failed: %s", cudaGetErrorString(cudaGetLastError()));
      float ms = 0.f;
      cudaEventElapsedTime(&ms, e0, e1);
      host_ms[k] = ms / iters;
    }
    if (exec) cudaGraphExecDestroy(exec);
  }
  pdl_regime() = saved_regime;
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  g_launches.store(before);
  return rc;
}

// Timeline of the last decode graph replay (only when the graph was captured with ECHR_TRACE=1):
// per kernel its kind (0 h2att, 1 attention, 2 gates, 3 logits, 4 finalize), earliest CTA start and
// latest CTA end in globaltimer ns.  Synchronises the device.  Returns the number of kernels.
extern "C" int echr_trace_read(echr_batch_t* b, int64_t* host_start_ns, int64_t* host_end_ns, int32_t* host_kind,
                               int64_t* host_detail, int max_entries) {
  if (!b || !b->trace_on || !host_start_ns || !host_end_ns || !host_kind) return 0;
  if (cudaDeviceSynchronize() != cudaSuccess) return 0;
  std::vector<unsigned long long> h(2 * TRACE_SLOTS);
  if (cudaMemcpy(h.data(), b->trace, sizeof(unsigned long long) * 2 * TRACE_SLOTS, cudaMemcpyDeviceToHost) != cudaSuccess)
    return 0;
  int n = b->trace_n < max_entries ? b->trace_n : max_entries;
  if (host_detail != nullptr &&
      cudaMemcpy(host_detail, b->trace_detail, sizeof(unsigned long long) * 4 * n, cudaMemcpyDeviceToHost) != cudaSuccess)
    return 0;
  for (int i = 0; i < n; ++i) {
    host_start_ns[i] = (int64_t)h[2 * i];
    host_end_ns[i] = (int64_t)(~h[2 * i + 1]);
    host_kind[i] = b->trace_kind[i];
  }
  return n;
}

extern "C" int64_t echr_launch_count(int reset) {
  const long long v = reset ? g_launches.exchange(0) : g_launches.load();
  return (int64_t)v;
}
