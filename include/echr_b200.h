/* echr_b200 — C-ABI of the B200-native (sm_100a) ECHR caption-generation hot path.
 *
 * The reference (ttengwang/ECHR) is pure Python/PyTorch and has no FFI of its own; the entry
 * points below are the ones a binding for its hot path would need.  Each one names
 * the reference code it replaces (paths relative to the reference repository root).
 *
 * Conventions
 *   - every function returns 0 on success or a negative code; echr_last_error() gives the text
 *     (thread-local).  Only the *_create calls allocate.
 *   - pointers are DEVICE pointers unless the parameter name starts with `host_` or the comment
 *     says host; sizes/strides are in elements.
 *   - `stream` is a cudaStream_t passed as void*; work is stream-ordered, nothing synchronises
 *     unless stated.
 *   - there is no CPU fallback: on a non-sm_100 device every entry fails with code -3.
 */
#ifndef ECHR_B200_H_
#define ECHR_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ECHR_B200_ABI_VERSION 5

#if defined(__GNUC__)
#define ECHR_API __attribute__((visibility("default")))
#else
#define ECHR_API
#endif

/* fixed sizes of the ECHR recipe (experiments/train_ECHR.sh:5, opts.py:81-154) */
#define ECHR_VIDEO_DIM 500  /* C3D clip feature */
#define ECHR_HIDDEN 512     /* every hidden / embedding size on the path */
#define ECHR_LDA_DIM 100
#define ECHR_HEADS 16
/* limits of one batch (checked by echr_batch_set_inputs, which fails with ECHR_ERR_INVALID beyond them):
 * events (proposals) per video — the reference's default selection feeds up to topN = 1000 (eval_utils.py:21,
 * 107-115); the relation encoder walks keys in shared-memory chunks, so the bound is the int32 pair count
 * (events^2 summed over the batch < 2^30), not shared memory;
 * clips per event span — the reference caps proposal spans at K = 256 clips (dataloader.py:515-520); the
 * attention kernels keep one score per clip of a window in shared memory. */
#define ECHR_MAX_EVENTS_PER_VIDEO 8192
#define ECHR_MAX_SPAN_CLIPS 1024

typedef struct echr_model echr_model_t;
typedef struct echr_batch echr_batch_t;

typedef struct echr_config {
  int32_t vocab_size; /* CG_vocab_size; the logit/embedding tables have vocab_size+1 rows        */
  int32_t seq_length; /* CG_seq_length (30)                                                       */
  int32_t a_terms;    /* 2: activations as bf16 hi+lo planes (fp32-grade); 1: one fp16 plane, and then   */
                      /*    the weights are fp16 too (one tcgen05 product per k-step)                 */
  int32_t w_terms;    /* 1: weights rounded to bf16 (fp16 when a_terms == 1), 2: bf16 hi+lo planes      */
  int32_t clip_bf16;  /* must be 0 (reserved).  The attention tables follow a_terms: fp32 (e^{2P}, fp32   */
                      /*    clips) with two planes, fp16 (P, fp16 clips) with one                        */
} echr_config_t;

/* fp32 device pointers into the reference checkpoint's `cg_model` state_dict (train.py:456-461;
 * key names in comments).  Dead parameters (h2a_layer, fusion_layer) are not needed. */
typedef struct echr_weights {
  const float* event_emb_w;  /* fusion_model.event_emb.weight            [512,1012] */
  const float* event_emb_b;  /*                       .bias              [512]      */
  const float* pos_fc1_w;    /* fusion_model.enc_attn.pair_pos_fc1.weight [512,512] */
  const float* pos_fc1_b;
  const float* pos_fc2_w;    /* fusion_model.enc_attn.pair_pos_fc2.weight [16,512]  */
  const float* pos_fc2_b;
  const float* query_w;      /* fusion_model.enc_attn.query_1.weight      [512,512] */
  const float* query_b;
  const float* key_w;        /* fusion_model.enc_attn.key_1.weight        [512,512] */
  const float* key_b;
  const float* out_w;        /* fusion_model.enc_attn.linear_out_1.weight [512,512,1,1] */
  const float* out_b;
  const float* embed_w;      /* lm_model.embed.weight                     [V+1,512] */
  const float* logit_w;      /* lm_model.logit.weight                     [V+1,1536] */
  const float* logit_b;
  const float* lstm_w_ih[3]; /* lm_model.core.layer{0,1,2}.weight_ih      [2048,1024|1012|612] */
  const float* lstm_w_hh[3]; /*                           .weight_hh      [2048,512] */
  const float* lstm_b_ih[3];
  const float* lstm_b_hh[3];
  const float* ctx2att_w;    /* lm_model.core.attention.ctx2att.weight    [512,500] */
  const float* ctx2att_b;
  const float* h2att_w;      /* lm_model.core.attention.h2att.weight      [512,512] */
  const float* h2att_b;
  const float* alpha_w;      /* lm_model.core.attention.alpha_net.weight  [1,512] (bias cancels in the softmax) */
} echr_weights_t;

typedef struct echr_capacity {
  int64_t max_videos;
  int64_t max_clips;  /* sum of clips over the videos of one batch   */
  int64_t max_events; /* sum of events (proposals) over the videos   */
  int64_t max_pairs;  /* sum over videos of events^2                 */
  int64_t max_rows;   /* decode rows = events * beam                 */
} echr_capacity_t;

/* One batch of independent videos.  All index arrays are HOST arrays. */
typedef struct echr_batch_desc {
  int32_t n_videos;
  int32_t n_events;
  const int32_t* host_clip_offsets; /* [n_videos+1] first clip row of each video in c3d/tap      */
  const int32_t* host_event_video;  /* [n_events]   video of each event (non-decreasing)          */
  const int32_t* host_event_start;  /* [n_events]   soi_select_list[i][0], video-local, inclusive */
  const int32_t* host_event_end;    /* [n_events]   soi_select_list[i][1], video-local, exclusive */
  const int32_t* host_event_ind;    /* [n_events]   ind_select_list[i],   video-local             */
} echr_batch_desc_t;

/* ---- library ---------------------------------------------------------------------------- */
ECHR_API const char* echr_last_error(void);
ECHR_API int echr_abi_version(void);
/* 0 if `device` is an sm_100 GPU this build can run on. */
ECHR_API int echr_device_check(int device);

/* ---- building blocks (also used by the unit parity tests) -------------------------------- */
/* fp32 [rows, cols] (row stride ld_src) -> bf16 planes hi (and lo = bf16(x - hi), may be NULL),
 * zero-padded to cols_dst columns, row stride ld_dst. */
ECHR_API int echr_split_bf16(const float* src, int64_t rows, int64_t cols, int64_t ld_src, void* hi, void* lo,
                    int64_t ld_dst, int64_t cols_dst, void* stream);
/* fp32 [rows, cols] -> ONE IEEE fp16 plane (saturated to +-65504), zero-padded like the above.  This is the
 * operand format of the single-plane products (a_planes == 1 in echr_gemm, a_terms == 1 in echr_config_t). */
ECHR_API int echr_split_f16(const float* src, int64_t rows, int64_t cols, int64_t ld_src, void* dst, int64_t ld_dst,
                   int64_t cols_dst, void* stream);

/* out[m, n] = act( A[m, k] * W[n, k]^T + bias[n] ) on the tcgen05 tensor cores.
 * Replaces every nn.Linear on the path (models/MA_attention_8_NEW.py:29,93-96;
 * models/OldModel_NEW.py:51,372-374).  A and W are 16-bit plane stacks ([planes][rows][ld]):
 * a_planes=2 is bf16 hi+lo (echr_split_bf16; W then bf16 with 1 or 2 planes, fp32-grade product);
 * a_planes=1 is ONE fp16 plane for A and for W (echr_split_f16; w_planes must be 1).
 * k must be a multiple of 64 (pad with zeros), k <= 2048.  act: 0 none, 1 tanh, 2 exp(2*clamp(x,+-40)).
 * Outputs: out_f32 (may be NULL) and/or planes out_hi/out_lo (out_hi alone = one fp16 plane, both = bf16 hi+lo).
 * block_n: 128, 64, 32 or 16 = one 128 x block_n tile per CTA; 256 = the persistent 128 x 256 kernel with
 * double-buffered TMEM accumulators the engine uses from 1024 rows up. */
ECHR_API int echr_gemm_planes(const void* a, int64_t m, int64_t k, int64_t a_ld, int a_planes,
                     int64_t a_plane_stride, const void* w, int64_t n, int64_t w_ld, int w_planes,
                     int64_t w_plane_stride, const float* bias, int act, float* out_f32,
                     int64_t ld_f32, void* out_hi, void* out_lo, int64_t ld_bf, int block_n,
                     void* stream);

/* One late-fusion attention step (models/OldModel_NEW.py:376-401 with ctx2att hoisted):
 * for each row r: window = clips [row_start[r], row_start[r]+row_len[r]) of proj/clip;
 * score_l = alpha . tanh(P[l] + q[r]); att[r] = sum_l softmax(score)_l * clip[l].
 * The projected operands come in the exponential domain the GEMM epilogues write (act = 2):
 * proj_exp [n_clips,512] = e^{2P} and q_exp [rows,512] = e^{2q}, exponents clamped to +-80; with them
 * tanh(P+q) = 1 - 2/(e^{2P} e^{2q} + 1) costs one reciprocal per element.
 * proj_exp and clip [n_clips, ld_clip] are fp32 (clip_bf16=0) or bf16 (clip_bf16=1); q_exp fp32;
 * out_f32 [rows,500] (may be NULL); out_hi/out_lo [rows, ld_bf] (may be NULL; out_hi alone = one fp16 plane,
 * both = bf16 hi+lo).
 * max_len: the largest row_len of the call if the caller knows it (lets small launches stage whole
 * windows in shared memory through the bulk-copy engine), 0 if unknown. */
ECHR_API int echr_attention_step(const float* q_exp, const void* proj_exp, const void* clip, int64_t ld_clip,
                        int clip_bf16, const float* alpha, const int32_t* row_start,
                        const int32_t* row_len, int64_t rows, float* out_f32, void* out_hi,
                        void* out_lo, int64_t ld_bf, int max_len, void* stream);

/* The same step over 16-bit tables (what the single-plane precision runs; csrc/attention_f16.cuh):
 * q_exp [rows,512] fp32 = e^{2 (h2att(h1) + b)}; tab16 [n_clips,1024] fp16, per clip
 * [ e^{2 (P - rowmax P)} (512) | clip features (500, zero padded to 512) ] with P = ctx2att(c3d) + b;
 * proj_scale [n_clips] fp32 = e^{2 rowmax P}.  One CTA per row; every warp streams its clips through a private
 * shared-memory ring filled by cp.async.bulk (one 2 KB copy per clip) and folds them into an online softmax; any
 * window length.  out_f32 [rows,500] (may be NULL); out16 [rows, ld_out] fp16 (may be NULL).  mean_len: the average
 * row_len if the caller knows it (launches of many short windows then run 2 or 4 warps per row instead of 8), 0 if
 * unknown. */
ECHR_API int echr_attention_step_f16(const float* q_exp, const void* tab16, const float* proj_scale, const float* alpha,
                            const int32_t* row_start, const int32_t* row_len, int64_t rows, float* out_f32,
                            void* out16, int64_t ld_out, int mean_len, void* stream);

/* ---- model: packed weights --------------------------------------------------------------- */
/* Packs the checkpoint into kernel layouts (bf16 planes, gate-interleaved LSTM rows, fused biases).
 * Replaces CaptionGenerator.__init__ + load_state_dict (CaptionGenerator.py:8-14, eval.py:65-67).
 * Synchronises `stream` before returning. */
ECHR_API int echr_model_create(const echr_config_t* config, const echr_weights_t* weights, void* stream,
                      echr_model_t** out);
ECHR_API int echr_model_destroy(echr_model_t* model);

/* ---- batch: workspace + inputs ----------------------------------------------------------- */
ECHR_API int echr_batch_create(const echr_model_t* model, const echr_capacity_t* capacity, echr_batch_t** out);
ECHR_API int echr_batch_destroy(echr_batch_t* batch);
/* c3d [n_clips,500], tap [n_clips,512], lda [n_videos,100] fp32; on_host=1: host buffers (pinned or
 * pageable), copied with cudaMemcpyAsync; on_host=0: device buffers.  Inputs are copied into the
 * batch workspace, so the caller's buffers may be reused once the stream has passed this call. */
ECHR_API int echr_batch_set_inputs(echr_batch_t* batch, const echr_batch_desc_t* desc, const float* c3d,
                          const float* tap, const float* lda, int on_host, void* stream);

/* Context assembly + event relation encoder for the whole batch: mean-pooled clips and proposal
 * states per event, pairwise position embedding, pair MLP, multi-head relation attention, and the
 * hoisted ctx2att projection (CaptionGenerator.py:87-167; models/MA_attention_8_NEW.py:35-177;
 * models/OldModel_NEW.py:381). */
ECHR_API int echr_encode(echr_batch_t* batch, void* stream);
/* event context [n_events,512] fp32 produced by echr_encode (CaptionGenerator.py:129). */
ECHR_API int echr_get_event_context(const echr_batch_t* batch, float* out, void* stream);

/* MA_Attention8.forward(feats, soi_select_list) on its own (models/MA_attention_8_NEW.py:35-49):
 * feats fp32 [n,1012] = (pooled clip feature | proposal state) per event of ONE video, host_soi int32
 * [n,2] half-open spans -> out fp32 [n,512].  Overwrites the batch's current inputs. */
ECHR_API int echr_relation_encode(echr_batch_t* batch, const float* feats, const int32_t* host_soi, int n, float* out,
                         void* stream);
/* lm_model-level contexts (models/OldModel_NEW.py:98,139 take ready-made video / event / clip):
 * video fp32 [100], event fp32 [n,512], clip fp32 [n,lmax,500] zero padded, host_clip_len int32 [n] valid
 * clips per event (the reference's clip_mask is a prefix mask).  Replaces set_inputs + encode. */
ECHR_API int echr_batch_set_context(echr_batch_t* batch, const float* video, const float* event, const float* clip,
                           const int32_t* host_clip_len, int n, int lmax, void* stream);
/* OldModel.get_logprobs_state (models/OldModel_NEW.py:133-137) with explicit state: it int32 [n],
 * h_in / c_in fp32 [3,n,512] -> logprobs fp32 [n,V+1], h_out / c_out fp32 [3,n,512]. */
ECHR_API int echr_decode_step(echr_batch_t* batch, const int32_t* it, const float* h_in, const float* c_in,
                     float* logprobs, float* h_out, float* c_out, void* stream);

/* Greedy decode resident on the device (models/OldModel_NEW.py:139-187, sample_max=1):
 * tokens int32 [n_events, seq_length] with the reference's masking (finished rows -> 0),
 * logp fp32 [n_events, seq_length] UNmasked max log-probabilities (as the reference returns). */
ECHR_API int echr_decode_greedy(echr_batch_t* batch, int32_t* tokens, float* logp, void* stream);
/* Beam search resident on the device (spec: oracle/echr_oracle.py beam_search; the reference has
 * none, models/OldModel_NEW.py:144-146).  beam in 1..4.  tokens/logp [n_events, seq_length] zero
 * padded, score [n_events] = cumulative log-probability of the returned hypothesis. */
ECHR_API int echr_decode_beam(echr_batch_t* batch, int beam, int32_t* tokens, float* logp, float* score,
                     void* stream);
/* The decisions of the last beam decode of this batch: for every step t and decode row r = event * beam + slot,
 * the parent slot (within the event) and the token the new slot was built from; parent / tok: int32 device
 * [seq_length, n_events * beam].  Lets a caller (the parity tests) replay the search with reference arithmetic. */
ECHR_API int echr_get_beam_trace(const echr_batch_t* batch, int beam, int32_t* parent, int32_t* tok, void* stream);
/* Teacher-forced log-probabilities (models/OldModel_NEW.py:98-130): seq int32 [n_events, seq_cols]
 * (column 0 = BOS = 0); logprobs fp32 [n_events, seq_cols-1, vocab_size+1]. */
ECHR_API int echr_teacher_forced(echr_batch_t* batch, const int32_t* seq, int seq_cols, float* logprobs,
                        void* stream);

/* Hint: `lanes` batches of this job run concurrently on the GPU (1 = this batch has it to itself, the
 * default).  With lanes > 1 the decode drops the choices that trade SM time for latency (cluster split-K of
 * the gate GEMMs, narrow h2att tiles).  Results do not change beyond fp32 summation order (tokens identical, |dlogp| < 1e-5). */
ECHR_API int echr_batch_set_concurrency(echr_batch_t* batch, int lanes);

/* The teacher-forced pass as the LM loss needs it (eval_utils.py:139-152; misc/utils.py:62-75
 * LanguageModelCriterion gathers one entry per step): target_logp fp32 [n_events, seq_cols-1],
 * target_logp[r, t] = log p(seq[r, t+1] | seq[r, 0..t]).  The [n, S-1, V+1] tensor is never written.
 * Allocates its scratch logit rows on first use (that call synchronises). */
ECHR_API int echr_teacher_forced_nll(echr_batch_t* batch, const int32_t* seq, int seq_cols, float* target_logp,
                            void* stream);

/* End-to-end call with HOST buffers (eval_utils.py:52,129-136 for a whole batch of videos):
 * H2D of the features, encode, decode (beam<=1: greedy), D2H of the results, stream synchronised
 * before returning.  host_score may be NULL. */
ECHR_API int echr_caption_host(echr_batch_t* batch, const echr_batch_desc_t* desc, const float* host_c3d,
                      const float* host_tap, const float* host_lda, int beam, int32_t* host_tokens,
                      float* host_logp, float* host_score, void* stream);
/* The same without the final synchronisation: the host buffers (pinned memory, or the copies are not
 * asynchronous) must stay valid and the outputs unread until `stream` has drained.  With two batches of one
 * model on two streams, the copies of one batch overlap the decode of the other. */
ECHR_API int echr_caption_host_async(echr_batch_t* batch, const echr_batch_desc_t* desc, const float* host_c3d,
                      const float* host_tap, const float* host_lda, int beam, int32_t* host_tokens,
                      float* host_logp, float* host_score, void* stream);

/* The same call with DEVICE buffers and nothing synchronised: inputs are copied into the workspace
 * (they may be reused as soon as the call returns in stream order), results are written to the device
 * arrays tokens/logp [n_events, seq_length] and score [n_events] (may be NULL).  One entry into the
 * library per batch: the host-side cost of a step is one description pass and two graph launches. */
ECHR_API int echr_caption_device(echr_batch_t* batch, const echr_batch_desc_t* desc, const float* c3d,
                        const float* tap, const float* lda, int beam, int32_t* tokens, float* logp,
                        float* score, void* stream);

/* ---- SST proposal encoder in front of the path -------------------------- */
/* models/sst_model.py:5-40: nn.LSTM(500 -> 512, 2 layers) over the clips of a video, then
 * sigmoid(Linear(512 -> K)).  fp32 DEVICE pointers into the tap_model state_dict. */
typedef struct echr_sst echr_sst_t;
typedef struct echr_sst_weights {
  const float* w_ih0;    /* rnn.weight_ih_l0 [2048, 500]  (gate order i,f,g,o as in nn.LSTM) */
  const float* w_hh0;    /* rnn.weight_hh_l0 [2048, 512] */
  const float* b_ih0;    /* rnn.bias_ih_l0   [2048]      */
  const float* b_hh0;    /* rnn.bias_hh_l0   [2048]      */
  const float* w_ih1;    /* rnn.weight_ih_l1 [2048, 512] */
  const float* w_hh1;    /* rnn.weight_hh_l1 [2048, 512] */
  const float* b_ih1;    /* rnn.bias_ih_l1   [2048]      */
  const float* b_hh1;    /* rnn.bias_hh_l1   [2048]      */
  const float* scores_w; /* scores.weight    [K, 512]    */
  const float* scores_b; /* scores.bias      [K]         */
} echr_sst_weights_t;
/* Packs the weights (a_terms / w_terms as in echr_config_t).  Synchronises `stream`. */
ECHR_API int echr_sst_create(int K, int a_terms, int w_terms, const echr_sst_weights_t* weights, void* stream,
                    echr_sst_t** out);
ECHR_API int echr_sst_destroy(echr_sst_t* sst);
/* SST.forward (sst_model.py:31-40, eval mode) for a batch of videos in one call: c3d fp32 [sum_v T_v, 500]
 * (device), host_clip_offsets int32 [n_videos+1] (host).  Outputs (device): tap_feats [sum_v T_v, 512] =
 * rnn_output, pred_proposals [sum_v T_v, K].  Videos of different length run together: the recurrent step t
 * covers the videos that still have a clip t.  The workspace grows on demand (that call synchronises). */
ECHR_API int echr_sst_forward(echr_sst_t* sst, const float* c3d, const int32_t* host_clip_offsets, int n_videos,
                     float* tap_feats, float* pred_proposals, void* stream);

/* ---- proposal selection between the proposal module and the path ------- */
/* All pointers are DEVICE pointers unless named host_*.  One CTA per video; the [T,K] score matrix stays in
 * HBM and only the selected records come out.  Outputs are fixed-capacity slots: video v writes its records
 * to [v*cap, v*cap + count[v]) of each output array, in the order the reference's lists have.
 *
 * gettop1000 (eval_utils.py:263-292).  scores, mask: fp32 [sum_v T_v, K] (mask may be NULL = all ones);
 * clip_offsets: int32 [n_videos+1].  Per video: p = scores*mask; thr = the topn-th largest p over all T_v*K
 * cells (the smallest when there are fewer); every cell (n,k) with n >= k and p >= max(thr, val_score_thres)
 * is emitted in (n,k) row-major order: out_ind = n (index_select_list), [out_start,out_end) = [n-k, n+1)
 * (featstamp_list), out_conf = p (confidence).  out_count[v] = number found; records past `cap` are dropped. */
ECHR_API int echr_select_topn(const float* scores, const float* mask, const int32_t* clip_offsets, int n_videos,
                     int K, int topn, float val_score_thres, int32_t* out_ind, int32_t* out_start,
                     int32_t* out_end, float* out_conf, int32_t* out_count, int cap, void* stream);

/* gettop1000_nms (eval_utils.py:295-331): greedy 1-D NMS over the proposals (n,k), n in [0,T), k in
 * [0,min(n,K)), [start,end) = [n-k, n+1), score = scores[n,k]: take the best remaining proposal, drop every
 * remaining one whose overlap (intersection+1)/(union) with it exceeds `overlap`, until none is left or topn
 * are taken.  Equal scores: the later proposal in enumeration order first (numpy's unstable argsort leaves
 * that case open).  Overlaps are evaluated in fp64 in numpy's operation order.
 * T_v and K at most 65535.  workspace: 8 bytes per proposal of the batch; workspace_offsets: int64 [n_videos] = first proposal of each
 * video (prefix sums of echr_select_nms_proposals(T_v, K)); total_proposals = their sum.
 * Outputs per pick, in pick order: out_ind = n (index_select_list), out_start/out_end (nms_props), out_score
 * (nms_scores), out_pick = the proposal's enumeration index (what indexes prop_gts). */
ECHR_API int64_t echr_select_nms_proposals(int T, int K);
ECHR_API int echr_select_nms(const float* scores, const int32_t* clip_offsets, int n_videos, int K, double overlap,
                    int topn, void* workspace, const int64_t* workspace_offsets, int64_t total_proposals,
                    int32_t* out_ind, int32_t* out_start, int32_t* out_end, float* out_score,
                    int32_t* out_pick, int32_t* out_count, int cap, void* stream);

/* gettopN_nms (eval_utils.py:230-260): fp64 t1, t2, prop_scores, sent_scores [sum_v n_v]; offsets int32
 * [n_videos+1].  While proposals remain (and fewer than topn are picked): i = best remaining by prop_score;
 * cluster = remaining with overlap(i,.) >= nms_overlap, overlap = (inter+1e-3)/(union); the pick is the cluster
 * member with the largest sent_score (first in ascending prop_score order on ties); remaining = those with
 * overlap(i,.) <= nms_overlap.  out_pick: video-local indices in pick order.  workspace: 1 byte per proposal. */
ECHR_API int echr_rerank_nms(const double* t1, const double* t2, const double* prop_scores,
                    const double* sent_scores, const int32_t* offsets, int n_videos, double nms_overlap,
                    int topn, void* workspace, int32_t* out_pick, int32_t* out_count, int cap, void* stream);

/* echr_rerank_nms without a workspace argument and without any allocation, for callers that cannot call
 * cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  It has NO
 * counterpart in the reference.  The one-byte-per-proposal
 * flags live in a static device array of ECHR_RERANK_STATIC_MAX bytes inside the library (held by every GPU that
 * loads it, until exit); total_proposals (= offsets[n_videos], which the caller knows on the host) beyond that
 * limit is refused with ECHR_ERR_INVALID.  Same inputs, outputs and results as echr_rerank_nms.  All calls on a
 * device share the array: calls that overlap on the GPU (different streams) race; calls ordered on one stream
 * do not, and nothing of a previous call's contents is read. */
#define ECHR_RERANK_STATIC_MAX (1 << 20)
ECHR_API int echr_rerank_nms_static(const double* t1, const double* t2, const double* prop_scores,
                    const double* sent_scores, const int32_t* offsets, int n_videos, int64_t total_proposals,
                    double nms_overlap, int topn, int32_t* out_pick, int32_t* out_count, int cap, void* stream);

/* Measurement aid for bench.py's roofline: average duration in ms of each kernel of one decode step
 * (host_ms[6] = h2att GEMM, attention, LSTM gate GEMM, logit GEMM, finalize, and the logit launch that also carries
 * the next step's attention query GEMM — what a greedy decode of >= 1024 rows runs instead of slots 3 + 0; 0 when
 * the batch's decode does not fuse them), from CUDA events on `stream` around a captured graph of `iters` back-to-back plain launches of that
 * kernel (no host launch gaps, as in the decode graph), with the schedule the batch's decode uses (see
 * echr_batch_set_concurrency).  Needs a decoded batch; synchronises. */
ECHR_API int echr_time_step_kernels(echr_batch_t* batch, int beam, int iters, float* host_ms, void* stream);

/* Debug aid: timeline of the last decode-graph replay when the graph was captured with ECHR_TRACE=1 in
 * the environment: per kernel its kind (0 h2att GEMM, 1 attention, 2 gate GEMM, 3 logit GEMM,
 * 4 finalize), earliest CTA start and latest CTA end (globaltimer ns); host_detail (may be NULL) gets 4
 * words per kernel, filled by the GEMMs only: sums over CTAs of (first operand stage landed, accumulator
 * ready, CTA end) minus CTA start in ns, and the CTA count.  Synchronises the device; returns the number
 * of entries written (0 when tracing is off). */
ECHR_API int echr_trace_read(echr_batch_t* batch, int64_t* host_start_ns, int64_t* host_end_ns, int32_t* host_kind,
                    int64_t* host_detail, int max_entries);

/* Number of kernel launches this library issued since the last call (graph replays count their
 * kernel nodes); used by bench.py's `gpu_launches`. */
ECHR_API int64_t echr_launch_count(int reset);

#ifdef __cplusplus
}
#endif
#endif /* ECHR_B200_H_ */
