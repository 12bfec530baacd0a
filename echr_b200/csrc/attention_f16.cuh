// Late-fusion attention step (models/OldModel_NEW.py:376-401, ctx2att hoisted) over 16-bit tables, staged through
// shared memory by the bulk-copy engine — the form BASELINE.json's north_star (2) asks for.
//
//   score_l = alpha . tanh(P_l + q)        P = ctx2att(c3d) + b
//   w       = softmax_l(score)             over the event's clip window [s0, s0 + L)
//   att     = sum_l w_l * clip_l           clip fp16 [clips,512] (500 used): plane 0 of the ctx2att operand
//
// Tables (written once per batch by the encoder).  tanh(x) = 1 - 2 / (e^{2x} + 1) and e^{2(P+q)} = e^{2P} e^{2q}, so
// with the projection stored in the exponential domain a score element costs ONE special-function operation (the
// reciprocal) instead of two.  e^{2P} does not fit fp16 as it is, so every clip row is stored relative to its own
// maximum: E_lk = e^{2 (P_lk - m_l)} in (0, 1] as fp16 plus S_l = e^{2 m_l} as fp32 [clips];
// e^{2(P+q)} = (E_lk S_l) e^{2 q_k}, the last factor coming from the h2att GEMM epilogue.  The fp16 rounding of E
// (2^-12 relative) moves tanh by at most 1.3e-4; entries more than 8 below the row maximum flush towards 0, where
// tanh = -1 to within e^{2(P+q)} itself.  No clamp is needed on the product: inf gives rcp -> 0, tanh = 1.
//
// Layout.  The table row and the fp16 clip features of a clip sit side by side, tab[clip] = [E (512) | clip (512)]
// (2 KB: the clip half is at the same time the A operand of the ctx2att GEMM, row stride 1024), so ONE bulk copy
// brings everything the step needs of a clip: the copy engine's cost is per request, not per byte, at this size.
//
// Schedule.  One CTA of W warps per decode row; warp w owns clips w, w+W, ... of the window and folds them into an
// online softmax (running max / sum / weighted feature sum, rescaled when the max moves); the W partial states are
// merged through shared memory at the end.  Every warp keeps a PRIVATE four-slot ring in shared memory that its
// lane 0 fills with cp.async.bulk (one slot = the 2 KB of one clip, one copy, one mbarrier) up to four clips ahead: 24 warps x 8 KB of HBM/L2 reads in flight per SM with no register held for
// them and no block-wide barrier inside a row.  Every byte of a window is read exactly once: no score buffer (so no
// bound on the window length), no second pass over the clip rows.
// W is chosen per launch: a row costs ~3 us of dependent latency (window descriptor -> first copies -> merge)
// whatever its length, so a launch with more rows than 2 x SMs runs 4 warps per row and 6 rows per SM (2 and 12 when
// the mean window is <= 16 clips); launches of few rows (the latency regime) or of unknown windows run 8 warps per row.
// Measured (us per launch, W = 2 / 4 / 8): 4072 rows, mean window 25: 40.7 / 41.0 / 52.8; 9600 rows, mean 50:
// - / 88 / 114; 3200 rows, mean 128: - / 95 / 96; 251 rows: 15.9 / 10.0 / 7.7.
// Algorithmic bytes per event-step: 2052 L + 2048 (e^{2q} in) + 1000 (att out).
// Other bounds: one reciprocal per (row, clip, hidden unit) on the special-function pipe (16 per clock and SM), 512 L
// per row-step: 30 us of the 84 us of a cfg3 launch, 31 of 91 (cfg4), 11 of 40 (cfg5); and the instruction stream,
// ~216 warp instructions per clip visit of which ~115 arithmetic: 55-60 % of the launch's issue slots.  Sharing a
// window's copies between the beam slots of an event changed the time by 3 %: long windows are not copy-bound.
#pragma once
#include "ptx.cuh"

namespace echr {

constexpr int ATR_SLOTS = 4;                              // clips in flight per warp
constexpr int ATR_SLOT_BYTES = 2048;                      // table row | clip row
constexpr int ATR_WARP_BYTES = ATR_SLOTS * ATR_SLOT_BYTES;   // per-warp ring (the rings double as the merge buffer)
constexpr int atr_smem(int warps) { return warps * ATR_WARP_BYTES + 512; }   // + mbarriers and merge scalars

struct AttF16Params {
  const float* q_exp;            // [rows, 512] fp32: e^{2 (h2att(h1) + b)}
  const __nv_bfloat16* tab;      // [clips, 1024] fp16 bits: E = e^{2 (P - rowmax P)} (512) | clip features (500, zero padded)
  const float* proj_scale;       // [clips] fp32: S = e^{2 rowmax P}
  const float* alpha;            // [512]
  const int32_t* row_start;      // [rows] first clip of the row's window
  const int32_t* row_len;        // [rows] clips in the window
  int rows;
  __nv_bfloat16* out;            // [rows, ld_out] fp16 bits (the ATT block of the activation matrix), may be null
  long long ld_out;
  float* out_f32;                // [rows, 500], may be null
};

}  // namespace echr
