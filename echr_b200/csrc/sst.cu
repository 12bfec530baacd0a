// SST proposal encoder in front of the caption path: the reference's
// models/sst_model.py:5-40 — nn.LSTM(500 -> 512, 2 layers, batch_first) over the T clips of a video followed by
// sigmoid(Linear(512 -> K)) — for a whole batch of ragged videos at once.
//
// Layout: videos are sorted by length (longest first) and everything recurrent lives in TIME-MAJOR matrices,
// row t*V + j = clip t of the j-th longest video, so the videos still running at step t are the first M_t rows
// of block t.  Per layer: one big tcgen05 GEMM computes the input projection of every clip (x W_ih^T + b_ih +
// b_hh, gate-interleaved columns) into a fp32 table, then T sequential GEMMs h_{t-1} W_hh^T with the LSTM
// epilogue of gemm_tc.cuh take their per-row bias from block t of that table and write h_t straight into block
// t+1 of the layer's operand planes (the next step's A operand and the next layer's input).  The reference runs
// cuDNN's LSTM one video per call (eval_utils.py:55).
#include "engine.cuh"
#include "util.cuh"

#include <algorithm>
#include <new>
#include <vector>

struct echr_sst {
  int K, a_terms, w_terms;
  void* arena;
  echr::PlaneMat w_ih[2], w_hh[2], w_score;
  float* b_gate[2];
  float* b_score;
  // workspace (grown on demand)
  void* ws;
  size_t ws_bytes;
  // The projection + recurrence + score launches of a batch shape (video count and the number of videos still
  // running at each step) are replayed from a CUDA graph: 2T+3 short dependent launches are host-bound otherwise.
  // A shape is run eagerly the first time it is seen and captured the second time; only workspace pointers are
  // baked into the graph (the gather / export kernels that touch the caller's buffers stay outside).
  struct Slot {
    unsigned long long key;
    cudaGraphExec_t exec;
    unsigned long long used;
  } slots[8];
  unsigned long long seen[16];
  int seen_next;
  unsigned long long clock;
  cudaStream_t cap_stream;  // capture happens here (the caller's stream may be the legacy default stream)
};

namespace echr {

// dst[4j+g, c] = w[g*512+j, c] for c < in_f, zero padded to in_pad: gate-interleaved rows (see engine.cu)
__global__ void sst_pack_gates_kernel(const float* __restrict__ w, int in_f, int in_pad, bf16* __restrict__ hi,
                                      bf16* __restrict__ lo, bool f16) {
  const long long total = (long long)GATES * in_pad;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i / in_pad), c = (int)(i - (long long)n * in_pad);
    const int j = n >> 2, g = n & 3;
    const float x = c < in_f ? w[(long long)(g * HID + j) * in_f + c] : 0.f;
    bf16 h, l;
    split16(x, f16, h, l);
    hi[i] = h;
    if (lo != nullptr) lo[i] = l;
  }
}
__global__ void sst_pack_bias_kernel(const float* __restrict__ b_ih, const float* __restrict__ b_hh,
                                     float* __restrict__ out) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n < GATES) {
    const int j = n >> 2, g = n & 3;
    out[n] = b_ih[g * HID + j] + b_hh[g * HID + j];
  }
}

// time-major operand planes of the clip features: row r = t*V + j <- c3d[src[r]] (zero row when src[r] < 0)
__global__ void __launch_bounds__(128)
sst_gather_clips_kernel(const float* __restrict__ c3d, const int32_t* __restrict__ src, bf16* __restrict__ hi,
                        bf16* __restrict__ lo) {
  const long long r = blockIdx.x;
  const int s = src[r];
  const int c0 = 4 * threadIdx.x;  // 128 threads x 4 columns = 512
  float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
  if (s >= 0 && c0 < ECHR_VIDEO_DIM) x = __ldg(reinterpret_cast<const float4*>(c3d + (long long)s * ECHR_VIDEO_DIM + c0));
  __align__(8) bf16 h[4], l[4];
  split16(x.x, lo == nullptr, h[0], l[0]); split16(x.y, lo == nullptr, h[1], l[1]);
  split16(x.z, lo == nullptr, h[2], l[2]); split16(x.w, lo == nullptr, h[3], l[3]);
  *reinterpret_cast<uint2*>(hi + r * CLIP_PAD + c0) = *reinterpret_cast<const uint2*>(h);
  if (lo != nullptr) *reinterpret_cast<uint2*>(lo + r * CLIP_PAD + c0) = *reinterpret_cast<const uint2*>(l);
}

// back to the caller's video-major order: tap_feats[clip] = top-layer h (from the operand planes),
// pred[clip] = the score row
__global__ void __launch_bounds__(128)
sst_export_kernel(const int32_t* __restrict__ src, const bf16* __restrict__ h_hi, const bf16* __restrict__ h_lo,
                  const float* __restrict__ scores, int K, float* __restrict__ tap, float* __restrict__ pred) {
  const long long r = blockIdx.x;
  const int s = src[r];
  if (s < 0) return;
  for (int c = threadIdx.x; c < HID; c += blockDim.x) {
    const float v = h_lo ? __bfloat162float(h_hi[r * HID + c]) + __bfloat162float(h_lo[r * HID + c])
                         : f16_value(h_hi[r * HID + c]);
    tap[(long long)s * HID + c] = v;
  }
  for (int c = threadIdx.x; c < K; c += blockDim.x) pred[(long long)s * K + c] = scores[r * K + c];
}

static cudaError_t sst_pack(const float* w, int in_f, const PlaneMat& dst, bool f16, cudaStream_t st) {
  sst_pack_gates_kernel<<<148 * 4, 256, 0, st>>>(w, in_f, (int)dst.ld, dst.plane(0),
                                                 dst.planes > 1 ? dst.plane(1) : nullptr, f16);
  count_launch();
  return cudaGetLastError();
}

static void sst_layout(echr_sst* m, Arena& a) {
  const int wt = m->w_terms;
  m->w_ih[0] = a.planes(wt, GATES, CLIP_PAD);
  m->w_ih[1] = a.planes(wt, GATES, HID);
  m->w_hh[0] = a.planes(wt, GATES, HID);
  m->w_hh[1] = a.planes(wt, GATES, HID);
  m->w_score = a.planes(wt, m->K, HID);
  m->b_gate[0] = a.take<float>(GATES);
  m->b_gate[1] = a.take<float>(GATES);
  m->b_score = a.take<float>(m->K);
}

struct SstWork {
  PlaneMat x, h[2];
  float* table;
  float* scores;
  float* c[2];
  int32_t* src;
};
static void sst_work_layout(const echr_sst* m, long long rows, int V, SstWork& w, Arena& a) {
  w.x = a.planes(m->a_terms, rows, CLIP_PAD);
  w.h[0] = a.planes(m->a_terms, rows + V, HID);
  w.h[1] = a.planes(m->a_terms, rows + V, HID);
  w.table = a.take<float>((size_t)rows * GATES);
  w.scores = a.take<float>((size_t)rows * m->K);
  w.c[0] = a.take<float>((size_t)V * HID);
  w.c[1] = a.take<float>((size_t)V * HID);
  w.src = a.take<int32_t>((size_t)rows);
}

}  // namespace echr

using namespace echr;

extern "C" int echr_sst_create(int K, int a_terms, int w_terms, const echr_sst_weights_t* w, void* stream,
                               echr_sst_t** out) {
  ECHR_REQUIRE(w && out, "echr_sst_create: null argument");
  ECHR_REQUIRE(K >= 1 && K <= 4096, "echr_sst_create: K out of range");
  ECHR_REQUIRE((a_terms == 1 || a_terms == 2) && (w_terms == 1 || w_terms == 2) && !(a_terms == 1 && w_terms == 2),
               "echr_sst_create: (a_terms, w_terms) must be (1,1), (2,1) or (2,2)");
  const float* const* wp = reinterpret_cast<const float* const*>(w);
  for (size_t i = 0; i < sizeof(echr_sst_weights_t) / sizeof(float*); ++i)
    ECHR_REQUIRE(wp[i] != nullptr, "echr_sst_create: weight pointer %d is null", (int)i);
  int rc = check_device();
  if (rc != ECHR_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  echr_sst* m = new (std::nothrow) echr_sst();
  ECHR_REQUIRE(m, "echr_sst_create: out of host memory");
  memset(m, 0, sizeof(*m));
  m->K = K; m->a_terms = a_terms; m->w_terms = w_terms;
  Arena sizing;
  sst_layout(m, sizing);
  if (cudaMalloc(&m->arena, sizing.off + 256) != cudaSuccess) {
    delete m;
    return set_error(ECHR_ERR_CUDA, "echr_sst_create: cudaMalloc(%zu) failed", sizing.off);
  }
  Arena a;
  a.base = reinterpret_cast<char*>(m->arena);
  sst_layout(m, a);
  const bool f16 = a_terms == 1;
  cudaError_t e = cudaMemsetAsync(m->arena, 0, sizing.off, st);
  if (e == cudaSuccess) e = sst_pack(w->w_ih0, ECHR_VIDEO_DIM, m->w_ih[0], f16, st);
  if (e == cudaSuccess) e = sst_pack(w->w_ih1, HID, m->w_ih[1], f16, st);
  if (e == cudaSuccess) e = sst_pack(w->w_hh0, HID, m->w_hh[0], f16, st);
  if (e == cudaSuccess) e = sst_pack(w->w_hh1, HID, m->w_hh[1], f16, st);
  if (e == cudaSuccess)
    e = launch_split(w->scores_w, K, HID, HID, m->w_score.plane(0), m->w_score.planes > 1 ? m->w_score.plane(1) : nullptr,
                     HID, HID, f16, st);
  if (e == cudaSuccess) {
    sst_pack_bias_kernel<<<GATES / 256, 256, 0, st>>>(w->b_ih0, w->b_hh0, m->b_gate[0]);
    sst_pack_bias_kernel<<<GATES / 256, 256, 0, st>>>(w->b_ih1, w->b_hh1, m->b_gate[1]);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(m->b_score, w->scores_b, sizeof(float) * K, cudaMemcpyDeviceToDevice, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) {
    cudaFree(m->arena);
    delete m;
    return set_error(ECHR_ERR_CUDA, "echr_sst_create: packing failed: %s", cudaGetErrorString(e));
  }
  *out = m;
  return ECHR_OK;
}

extern "C" int echr_sst_destroy(echr_sst_t* m) {
  if (!m) return ECHR_OK;
  for (auto& sl : m->slots)
    if (sl.exec != nullptr) cudaGraphExecDestroy(sl.exec);
  if (m->cap_stream != nullptr) cudaStreamDestroy(m->cap_stream);
  cudaFree(m->ws);
  cudaFree(m->arena);
  delete m;
  return ECHR_OK;
}

extern "C" int echr_sst_forward(echr_sst_t* m, const float* c3d, const int32_t* host_clip_offsets, int n_videos,
                                float* tap_feats, float* pred_proposals, void* stream) {
  ECHR_REQUIRE(m && c3d && host_clip_offsets && tap_feats && pred_proposals, "echr_sst_forward: null argument");
  ECHR_REQUIRE(n_videos >= 0, "echr_sst_forward: bad video count");
  cudaStream_t st = (cudaStream_t)stream;
  const int V = n_videos;
  if (V == 0) return ECHR_OK;
  // longest first
  std::vector<int> order(V), len(V);
  int tmax = 0;
  for (int v = 0; v < V; ++v) {
    len[v] = host_clip_offsets[v + 1] - host_clip_offsets[v];
    ECHR_REQUIRE(len[v] >= 0, "echr_sst_forward: clip offsets must not decrease");
    order[v] = v;
    tmax = std::max(tmax, len[v]);
  }
  if (tmax == 0) return ECHR_OK;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return len[a] > len[b]; });
  const long long rows = (long long)tmax * V;
  ECHR_REQUIRE(rows < (1LL << 31), "echr_sst_forward: batch too large");
  std::vector<int32_t> src((size_t)rows, -1);
  std::vector<int> active(tmax, 0);
  for (int j = 0; j < V; ++j) {
    const int v = order[j];
    for (int t = 0; t < len[v]; ++t) {
      src[(size_t)t * V + j] = host_clip_offsets[v] + t;
      active[t] = j + 1;
    }
  }
  Arena sizing;
  SstWork w;
  sst_work_layout(m, rows, V, w, sizing);
  if (sizing.off + 256 > m->ws_bytes) {
    ECHR_CUDA(cudaStreamSynchronize(st));
    cudaFree(m->ws);
    m->ws = nullptr;
    m->ws_bytes = 0;
    for (auto& sl : m->slots)  // the graphs hold pointers into the old workspace
      if (sl.exec != nullptr) { cudaGraphExecDestroy(sl.exec); sl.exec = nullptr; }
    const size_t want = sizing.off + 256 + (sizing.off >> 2);
    if (cudaMalloc(&m->ws, want) != cudaSuccess)
      return set_error(ECHR_ERR_WORKSPACE, "echr_sst_forward: cudaMalloc(%zu) for the workspace failed", want);
    m->ws_bytes = want;
  }
  Arena a;
  a.base = reinterpret_cast<char*>(m->ws);
  sst_work_layout(m, rows, V, w, a);
  const int at = m->a_terms, wt = m->w_terms;
  // the index table comes from pageable host memory: the copy returns once it is staged
  ECHR_CUDA(cudaMemcpyAsync(w.src, src.data(), sizeof(int32_t) * (size_t)rows, cudaMemcpyHostToDevice, st));
  sst_gather_clips_kernel<<<(unsigned)rows, 128, 0, st>>>(c3d, w.src, w.x.plane(0), at > 1 ? w.x.plane(1) : nullptr);
  count_launch();
  ECHR_CUDA(cudaGetLastError());
  auto enqueue = [&](cudaStream_t st) -> int {
    for (int l = 0; l < 2; ++l) {
      // input projection of every clip of the layer
      const PlaneMat in = l == 0 ? w.x : w.h[0].row_view(V, rows);
      GemmSpec p;
      p.A = &in; p.M = (int)rows; p.W = &m->w_ih[l]; p.N = GATES; p.K = l == 0 ? CLIP_PAD : HID;
      p.bias = m->b_gate[l];
      p.out_f32 = w.table; p.ld_f32 = GATES;
      int rc = run_gemm(&p, 1, at, wt, 128, EPI_STORE, 1, st, "sst input projection");
      if (rc != ECHR_OK) return rc;
      // h_{-1} = 0, c_{-1} = 0
      for (int pl = 0; pl < at; ++pl)
        ECHR_CUDA(cudaMemsetAsync(w.h[l].plane(pl), 0, sizeof(bf16) * (size_t)V * HID, st));
      ECHR_CUDA(cudaMemsetAsync(w.c[0], 0, sizeof(float) * (size_t)V * HID, st));
      for (int t = 0; t < tmax; ++t) {
        const PlaneMat hin = w.h[l].row_view((long long)t * V, V);
        GemmSpec s;
        s.A = &hin; s.M = active[t]; s.W = &m->w_hh[l]; s.N = GATES; s.K = HID;
        s.row_bias = w.table + (size_t)t * V * GATES; s.ld_row_bias = GATES;
        s.c_in = w.c[t & 1]; s.c_out = w.c[(t + 1) & 1];
        s.out_hi = w.h[l].plane(0) + (size_t)(t + 1) * V * HID;
        s.out_lo = at > 1 ? w.h[l].plane(1) + (size_t)(t + 1) * V * HID : nullptr;
        s.ld_bf = HID;
        rc = run_gemm(&s, 1, at, wt, 128, EPI_LSTM, 1, st, "sst recurrent step");
        if (rc != ECHR_OK) return rc;
      }
    }
    {
      const PlaneMat top = w.h[1].row_view(V, rows);
      GemmSpec p;
      p.A = &top; p.M = (int)rows; p.W = &m->w_score; p.N = m->K; p.K = HID;
      p.bias = m->b_score; p.act = ACT_SIGMOID;
      p.out_f32 = w.scores; p.ld_f32 = m->K;
      const int rc = run_gemm(&p, 1, at, wt, 128, EPI_STORE, 1, st, "sst scores");
      if (rc != ECHR_OK) return rc;
    }
    return ECHR_OK;
  };
  // shape key: FNV-1a over (V, tmax, active[])
  unsigned long long key = 1469598103934665603ull;
  auto mix = [&](unsigned long long x) { key = (key ^ x) * 1099511628211ull; };
  mix((unsigned long long)V); mix((unsigned long long)tmax);
  for (int t = 0; t < tmax; ++t) mix((unsigned long long)active[t]);
  key |= 1ull;
  echr_sst::Slot* hit = nullptr;
  for (auto& sl : m->slots) if (sl.exec != nullptr && sl.key == key) hit = &sl;
  static const bool use_graph = !(knob("ECHR_SST_GRAPH") && atoi(knob("ECHR_SST_GRAPH")) == 0);
  if (hit == nullptr && use_graph) {
    bool seen_before = false;
    for (unsigned long long k : m->seen) seen_before = seen_before || k == key;
    if (!seen_before) {
      m->seen[m->seen_next] = key;
      m->seen_next = (m->seen_next + 1) % 16;
    } else {
      echr_sst::Slot* victim = &m->slots[0];
      for (auto& sl : m->slots) if (sl.exec == nullptr || sl.used < victim->used) { victim = &sl; if (sl.exec == nullptr) break; }
      if (victim->exec != nullptr) { cudaGraphExecDestroy(victim->exec); victim->exec = nullptr; }
      cudaGraph_t graph = nullptr;
      if (m->cap_stream == nullptr) ECHR_CUDA(cudaStreamCreateWithFlags(&m->cap_stream, cudaStreamNonBlocking));
      ECHR_CUDA(cudaStreamBeginCapture(m->cap_stream, cudaStreamCaptureModeRelaxed));
      const int rc = enqueue(m->cap_stream);
      const cudaError_t ce = cudaStreamEndCapture(m->cap_stream, &graph);
      if (rc != ECHR_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
      if (ce != cudaSuccess) return set_error(ECHR_ERR_CUDA, "echr_sst_forward: capture failed: %s", cudaGetErrorString(ce));
      const cudaError_t ie = cudaGraphInstantiate(&victim->exec, graph, 0);
      cudaGraphDestroy(graph);
      if (ie != cudaSuccess) { victim->exec = nullptr; return set_error(ECHR_ERR_CUDA, "echr_sst_forward: instantiate failed: %s", cudaGetErrorString(ie)); }
      victim->key = key;
      hit = victim;
    }
  }
  if (hit != nullptr) {
    hit->used = ++m->clock;
    ECHR_CUDA(cudaGraphLaunch(hit->exec, st));
    count_launch(2 * tmax + 3);
  } else {
    const int rc = enqueue(st);
    if (rc != ECHR_OK) return rc;
  }
  sst_export_kernel<<<(unsigned)rows, 128, 0, st>>>(w.src, w.h[1].plane(0) + (size_t)V * HID,
                                                   at > 1 ? w.h[1].plane(1) + (size_t)V * HID : nullptr, w.scores, m->K,
                                                   tap_feats, pred_proposals);
  count_launch();
  ECHR_CUDA(cudaGetLastError());
  // `src` (host vector) must outlive the staged copy: pageable cudaMemcpyAsync returns after staging
  return ECHR_OK;
}
