// Late-fusion attention step (models/OldModel_NEW.py:376-401, ctx2att hoisted) over 16-bit tables, staged through
// shared memory by the bulk-copy engine — the form BASELINE.json's north_star (2) asks for.
//
//   score_l = alpha . tanh(P_l + q)        P = ctx2att(c3d) + b   fp16 [clips,512], written once per batch
//   w       = softmax_l(score)             over the event's clip window [s0, s0 + L)
//   att     = sum_l w_l * clip_l           clip fp16 [clips,512] (500 used): plane 0 of the ctx2att operand
//
// Persistent CTAs (8 warps, two per SM) walk the decode rows round robin.  A row's window is contiguous in HBM, so
// it is streamed in chunks of 8 clips by cp.async.bulk — two copies per chunk (8 projected rows, 8 clip rows:
// 16 KB) on one mbarrier — through a four-stage ring that thread 0 keeps full ACROSS row boundaries (the first
// chunks of the next row are in flight while the current row is being finished): ~96 KB of HBM/L2 reads in flight
// per SM without a register held for them, which is what a 6.5 TB/s stream at ~2 us loaded latency needs.  Warp w
// consumes clip w of every chunk, so a row's clips are processed eight at a time.  Each warp folds its clips into
// an online softmax (running max / sum / weighted feature sum, rescaled when the max moves); the eight partial
// states are merged through shared memory at the end of the row.  Every byte of a window is read exactly once: no
// score buffer, no second pass over the clip rows.  Algorithmic bytes per event-step: 2048 L + 2048 (q in) + 1000
// (att out).
//
// tanh(P + q) = 1 - 2 / (e^{2(P+q)} + 1): one FFMA + EX2 + FADD + RCP + FFMA per element, all in fp32; the
// constant sum_k alpha_k and the alpha_net bias drop out of the softmax.  No clamp is needed: e^{2x} -> inf gives
// rcp -> 0, tanh = 1.  The only 16-bit roundings are those of the stored tables (fp16: 2^-12 relative).
#pragma once
#include "ptx.cuh"

namespace echr {

constexpr int ATR_WARPS = 8;                             // = clips per chunk
constexpr int ATR_STAGES = 4;
constexpr int ATR_STAGE_BYTES = ATR_WARPS * 2048;        // 8 projected rows | 8 clip rows
constexpr int ATR_MAX_ROWS_PER_CTA = 256;                // row table in shared memory
// ring | merge buffer [8][512] f32 | alpha | row table (first clip, length) | barriers + merge scalars
constexpr int ATR_SMEM = ATR_STAGES * ATR_STAGE_BYTES + ATR_WARPS * 2048 + 2048 + ATR_MAX_ROWS_PER_CTA * 8 + 128;

struct AttF16Params {
  const float* q;                // [rows, 512] fp32: h2att(h1) + b  (NOT exponentiated)
  const __nv_bfloat16* proj;     // [clips, 512] fp16 bits
  const __nv_bfloat16* clip;     // [clips, 512] fp16 bits, columns 500..511 zero
  const float* alpha;            // [512]
  const int32_t* row_start;      // [rows] first clip of the row's window
  const int32_t* row_len;        // [rows] clips in the window
  int rows;
  __nv_bfloat16* out;            // [rows, ld_out] fp16 bits (the ATT block of the activation matrix), may be null
  long long ld_out;
  float* out_f32;                // [rows, 500], may be null
};

}  // namespace echr
