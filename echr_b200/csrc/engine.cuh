// Internal data model of the engine: packed weights (Model) and the per-batch workspace (Batch).
// Layout notes are in DESIGN.md §"Data layout in HBM".
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#include "../../include/echr_b200.h"
#include "gemm_tc.cuh"

namespace echr {

typedef __nv_bfloat16 bf16;

// ---- activation matrix: one row per decode row, bf16, `planes` planes, two ping-pong buffers.
// Column blocks of 64 (= one TMA box / one GEMM k-block).
constexpr int ACT_X = 0;       // embedded input token      512  (kb 0..7)
constexpr int ACT_EV = 512;    // event context (relation)  512  (kb 8..15)
constexpr int ACT_ATT = 1024;  // attended clip feature     500 -> 512 (kb 16..23)
constexpr int ACT_VID = 1536;  // video (LDA) context       100 -> 128 (kb 24..25)
constexpr int ACT_H = 1664;    // h of stream 0,1,2         3 x 512    (kb 26..49)
constexpr int ACT_LD = 3200;
constexpr int HID = 512;
constexpr int VID_PAD = 128;
constexpr int CLIP_PAD = 512;   // 500 -> 512
constexpr int FEAT_PAD = 1024;  // 1012 -> 1024
constexpr int GATES = 2048;

struct PlaneMat {  // [planes][rows][ld] 16-bit (bf16 hi/lo planes, or one fp16 plane)
  bf16* p = nullptr;
  int64_t rows = 0, ld = 0, planes = 0;
  int64_t stride = 0;  // elements between planes; 0 = rows * ld (a row-range view keeps its parent's stride)
  int64_t plane_stride() const { return stride ? stride : rows * ld; }
  PlaneMat row_view(int64_t row0, int64_t n) const {
    PlaneMat v = *this;
    v.p = p + row0 * ld;
    v.rows = n;
    v.stride = plane_stride();
    return v;
  }
  bf16* plane(int i) const { return p + i * plane_stride(); }
};

}  // namespace echr

struct echr_model {
  echr_config_t cfg;
  int V1;  // vocab_size + 1
  // tensor-core weights as bf16 planes (w_terms planes), row-major [out, in_padded]
  echr::PlaneMat w_event_emb;  // [512, 1024]
  echr::PlaneMat w_qky;        // [1536, 512]  rows: query | key | linear_out (grouped conv as dense)
  echr::PlaneMat w_fc1;        // [512, 512]
  echr::PlaneMat w_fc2;        // [16, 512]
  echr::PlaneMat w_ctx2att;    // [512, 512(500)]
  echr::PlaneMat w_h2att;      // [512, 512]
  echr::PlaneMat w_lstm[3];    // [2048 gate-interleaved, K_l]  K = 1536, 1536, 1152 (x | h | ctx)
  echr::PlaneMat w_logit;      // [V1, 1536]
  echr::PlaneMat embed;        // [V1, 512] always 2 planes (a_terms are read from it)
  // fp32 vectors
  float *b_event_emb, *b_qky, *b_fc1, *b_fc2, *b_ctx2att, *b_h2att, *b_lstm[3], *b_logit, *b_out,
      *alpha;
  // token -> gate pre-activation tables, [V1, 2048] fp32 per stream, gate-interleaved columns:
  // tok_gates[l][v] = W_ih_l[:, 0:512] . embed[v]  (+ b_ih + b_hh for stream 1, whose other input terms
  // all go through the step GEMM).  The embedding lookup and the x block of every gate GEMM become one
  // row gather in the LSTM epilogue.
  float* tok_gates[3];
  void* arena;  // one allocation behind everything above
};

struct echr_batch {
  const echr_model* model;
  echr_capacity_t cap;
  // current batch
  int n_videos, n_events, n_clips, n_pairs, max_n;  // max_n = largest events-per-video
  int max_len;                                      // longest event span (clips)
  int mean_len;                                     // mean event span (clips), rounded up
  bool encoded;
  // inputs (copies)
  float *c3d, *tap, *lda;  // [clips,500]; tap_feats[ind] per event [events,512]; [videos,100]
  float* host_tap_stage;   // pinned [max_events,512]: host-side gather of the proposal states
  // both pinned staging areas are double-buffered: a call fills the half the previous call did not use and
  // only waits (stage_ev) for the copy that read that half two calls ago
  float* host_tap_base;
  int32_t* host_stage_base;
  cudaEvent_t stage_ev[2];
  int stage_idx;
  // per-event int arrays on device: video, first clip (global), length, ind (global), pair row
  // offset, first event of the video, events in the video
  int32_t *ev_video, *ev_start, *ev_len, *ev_ind, *ev_pair_off, *ev_first, *ev_n;
  int32_t *vid_first, *vid_n;  // per video: first event, number of events
  int32_t* host_stage;  // pinned staging for the arrays above
  // encode workspace
  echr::PlaneMat clip_pl;   // [clips, 512]   c3d planes (A of ctx2att).  Single-plane precision: a view into the
                            // attention table [clips, 1024] = [e^{2 (P - rowmax)} | clip], row stride 1024
  float* proj_f32;          // [clips, 512]   two-plane precisions: e^{2P}, P = ctx2att(c3d)+b; single-plane: P (scratch)
  echr::bf16* proj_h;       // [clips, 1024]  single-plane precision: the attention table (attention_f16.cuh)
  float* proj_scale;        // [clips]        single-plane precision: e^{2 rowmax P}
  echr::PlaneMat feat_pl;   // [events, 1024] (mean c3d | tap[ind])
  echr::PlaneMat x_pl;      // [events, 512]  event_emb output
  float* qky;               // [events, 1536]
  float* pair_w;            // [16][n_pairs]  pair weights, head-major (pair_mlp.cu)
  float* event_ctx;         // [events, 512]
  // decode workspace (rows = events * beam)
  echr::PlaneMat act[2];    // [rows, ACT_LD]
  float* cstate[2];         // [3][rows][512]
  float* q_att;             // [rows, 512]   e^{2q}
  float* row_bias[2];       // [rows, 2048]  hoisted gate bias of streams 0 and 2
  float* partial;           // [rows, n_vtiles, 12]
  int32_t *row_start, *row_len;  // attention window per decode row
  int32_t* unfinished;      // [rows]
  int32_t* row_tok;         // [rows] input token of the next step (BOS = 0 at step 0)
  int32_t* row_parent;      // [rows] beam search: row whose cell state this row continues at the next step
  unsigned long long* argmax_key[2];  // [rows * KEY_STRIDE] packed running arg-max of the logits (greedy), by step parity
  int32_t* tokens_dev;      // [events, S] scratch outputs for echr_caption_host
  float* logp_dev;
  float* score_dev;
  // beam state
  float* beam_score;        // [rows]
  int32_t* beam_live;       // [rows]
  int32_t* beam_hist_tok;   // [2][rows][S]
  float* beam_hist_lp;      // [2][rows][S]
  float* best_score;        // [events]
  int32_t* beam_trace;      // [2][S][max_rows]: parent slot / token every new slot was built from, per step
  // CUDA graph cache (LRU): decode graphs keyed by (rows, beam mode, window-length bucket), encode graphs
  // keyed by the batch shape.  A key holds every host-side value the enqueue code bakes into the launches.
  struct GraphSlot {
    long long key[6];
    cudaGraphExec_t exec;
    long long kernels;  // kernel nodes in the graph
    unsigned long long stamp;
  };
  // > 1: other batches of the job run concurrently on this GPU (CaptionPipeline lanes).  Choices that buy latency
  // with extra SM time (cluster split-K of the gate GEMMs: two CTAs and a DSMEM exchange per tile; narrow h2att
  // tiles) are then off: with the SMs shared by several decode chains, SM time is the budget (cfg2, 4 lanes:
  // 329k -> 381k events/s).
  int concurrency;
  // scratch rows of raw logits for the teacher-forced loss path (allocated on its first use)
  float* nll_logits;
  long long nll_rows;
  static constexpr int GRAPH_SLOTS = 64;
  GraphSlot graphs[GRAPH_SLOTS];
  unsigned long long graph_clock;
  // shapes seen once are run eagerly: a shape is only captured when it comes back (a per-video eval loop sees a
  // new clip count on almost every call, and capturing + instantiating a 150-node graph costs ~20 ms)
  static constexpr int SEEN_KEYS = 64;
  long long seen_keys[SEEN_KEYS][6];
  int seen_next;
  cudaStream_t cap_stream;   // private stream graphs are captured on
  cudaStream_t cap_stream2;  // forked branches inside a capture
  cudaStream_t cap_stream3;
  cudaEvent_t cap_events[4]; // fork / join markers, reused inside a capture
  // timeline trace (ECHR_TRACE=1): device [2*TRACE_SLOTS] u64, host-side kind of every slot
  unsigned long long* trace;
  unsigned long long* trace_detail;  // [TRACE_SLOTS][4], GEMM kernels only (see gemm_tc.cuh)
  int trace_on, trace_n;
  unsigned char trace_kind[1024];
  void* arena;
  size_t arena_bytes;
};

namespace echr {

// ---- kernels_ctx.cu
cudaError_t launch_split(const float* src, int64_t rows, int64_t cols, int64_t ld_src, bf16* hi,
                         bf16* lo, int64_t ld_dst, int64_t cols_dst, bool f16, cudaStream_t st);
cudaError_t launch_event_feats(const echr_batch* b, cudaStream_t st);
cudaError_t launch_gather_tap(const echr_batch* b, const float* tap_src, cudaStream_t st);
cudaError_t launch_relation_attn(const echr_batch* b, cudaStream_t st);
// ---- pair_mlp.cu: position embedding -> fc1 + tanh -> fc2 fused over tiles of 128 event pairs
int launch_pair_mlp(const echr_batch* b, cudaStream_t st);
// ---- kernels_step.cu
cudaError_t launch_attention(const float* q, const void* proj, const void* clip, int64_t ld_clip,
                             int clip_bf16, const float* alpha, const int32_t* row_start,
                             const int32_t* row_len, int64_t rows, float* out_f32, bf16* out_hi,
                             bf16* out_lo, int64_t ld_bf, cudaStream_t st, int max_len = 0,
                             unsigned long long* trace = nullptr, int trace_slot = 0);
cudaError_t launch_attention_f16(const AttF16Params& P, int mean_len, cudaStream_t st, unsigned long long* trace = nullptr, int trace_slot = 0);
cudaError_t launch_attention_table(const float* proj, int clips, bf16* table, long long ld, float* scale, cudaStream_t st);
cudaError_t launch_decode_init(const echr_batch* b, int beam, cudaStream_t st);
cudaError_t launch_finalize_greedy(echr_batch* b, int step, int32_t* tokens, float* logp,
                                   cudaStream_t st, bool keys = false);
cudaError_t launch_finalize_beam(const echr_batch* b, int beam, int step, int32_t* tokens,
                                 float* logp, float* score, cudaStream_t st);
cudaError_t launch_finalize_tf(const echr_batch* b, int step, const int32_t* seq, int seq_cols,
                               float* logprobs, cudaStream_t st);
cudaError_t launch_finalize_nll(const echr_batch* b, int step, const int32_t* seq, int seq_cols,
                                const float* logits, float* target_logp, cudaStream_t st);
cudaError_t launch_load_state(const echr_batch* b, int rows, const int32_t* it, const float* h_in, cudaStream_t st);
cudaError_t launch_export_state(const echr_batch* b, int rows, float* h_out, cudaStream_t st);
void count_launch(int n = 1);

// ---------------------------------------------------------------- arena helper
struct Arena {
  size_t off = 0;
  char* base = nullptr;
  template <typename T>
  T* take(size_t n) {
    off = (off + 255) & ~(size_t)255;
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += n * sizeof(T);
    return p;
  }
  PlaneMat planes(int64_t planes_, int64_t rows, int64_t ld) {
    PlaneMat m;
    m.rows = rows; m.ld = ld; m.planes = planes_;
    m.p = take<bf16>((size_t)planes_ * rows * ld);
    return m;
  }
};


// ---------------------------------------------------------------- GEMM problem helper
struct GemmSpec {
  const PlaneMat* A = nullptr;
  int a_col0 = 0;        // first activation column (multiple of 64) when the k-blocks are contiguous
  const int* a_kb = nullptr;  // or an explicit k-block table
  int M = 0;
  const PlaneMat* W = nullptr;
  int N = 0, K = 0;
  const float* bias = nullptr;
  int act = 0;  // ACT_NONE
  bool epi_store = false;  // inside a persistent EPI_LOGITS launch: store epilogue for this problem (gemm_tc2.cuh)
  float* out_f32 = nullptr;
  long long ld_f32 = 0;
  bf16* out_hi = nullptr;
  bf16* out_lo = nullptr;
  long long ld_bf = 0;
  // LSTM
  const float* c_in = nullptr;
  const int32_t* c_src = nullptr;  // see GemmProblem::c_src
  float* c_out = nullptr;
  const float* row_bias = nullptr;
  long long ld_row_bias = 0;
  int hi_only_first = 0;  // leading k-blocks whose low-order activation plane is identically zero
  const int32_t* row_tok = nullptr;
  const float* tok_table = nullptr;
  long long ld_tok_table = 0;
  int ksplit = 1;  // 2: cluster split-K (run_gemm then launches the SPLIT == 2 kernel for the whole group)
  const unsigned long long* row_key = nullptr;  // token from the packed arg-max keys instead of row_tok
  unsigned long long* argmax_key = nullptr;     // logits: running arg-max keys
  // logits
  float* partial = nullptr;
  float* logits_out = nullptr;
  long long ld_logits = 0;
};


// One launch of `groups` independent GEMM problems (see gemm_tc.cuh); epi = EPI_STORE / EPI_LSTM / EPI_LOGITS.
int run_gemm(const GemmSpec* specs, int groups, int a_terms, int w_terms, int block_n, int epi, int topk,
             cudaStream_t st, const char* what, echr_batch* tb = nullptr, int kind = 5);
int check_device();
// partial records per row the logit GEMM writes when it runs over `rows` decode rows (depends on the
// GEMM variant the engine picks for that row count)
int logit_groups_for(const echr_model* m, int rows);
constexpr int TRACE_SLOTS = 1024;
// next trace slot of this batch for a kernel of `kind` (0 h2att, 1 attention, 2 gates, 3 logits, 4 finalize,
// 5 other); returns -1 / nullptr when tracing is off
int trace_next(echr_batch* b, int kind);
inline unsigned long long* trace_ptr(const echr_batch* b) { return b->trace_on == 1 ? b->trace : nullptr; }

}  // namespace echr
