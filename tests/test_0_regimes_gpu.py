"""GPU parity tests at the shapes and schedules every benchmark number rests on (collected first: `test_0_`).

Each test drives the C-ABI path and compares it with oracle/echr_oracle.py (pinned to the reference, see
tests/test_oracle_vs_reference.py and tests/golden/) on the same seeded weights and inputs:
  * cfg2 whole (BASELINE configs[1]: 64 videos / 251 events) in the benched precision (f16) through a
    CaptionPipeline of depth 4 — the shared-GPU schedule of bench.py — from device and from host buffers;
  * >= 1024 decode rows: the persistent 128 x 256 GEMM with its LSTM / logit epilogues and the streaming attention;
  * one cfg3 video (100 SST-style proposals, beam 3) and the whole cfg3 batch (9600 beam rows);
  * one cfg4 video (400 clips, 200 proposals, spans up to 256 clips).
Tolerances are BASELINE.json's: tokens identical except where the oracle's top-2 (greedy) / kept-vs-dropped (beam)
score gap is below 1e-3 (beam: every decision of the search is replayed with the oracle's arithmetic, see
helpers.check_beam_replay); per-step log-probs within 1e-4 (fp32-grade) or 2e-2 (16-bit operands)."""
import numpy as np
import pytest
import torch

import helpers
from echr_b200 import synth
from echr_b200.caption_generator import CaptionGenerator
from echr_b200.config import echr_opt
from echr_b200.engine import CaptionPipeline
from oracle import echr_oracle as O

pytestmark = pytest.mark.gpu
torch.set_grad_enabled(False)
DEV = "cuda:0"
# per-step log-prob tolerance (north_star) and the bound this implementation actually holds
LOGP_ALLOWED = {"fp32": 1e-4, "bf16w": 1e-4, "f16": 2e-2}
LOGP_HELD = {"fp32": 1e-4, "bf16w": 1e-4, "f16": 2e-3}
_models = {}


def model(precision):
    """fp32 precision on the fp32 checkpoint; bf16w / f16 on the bf16-rounded checkpoint (both sides of a 16-bit
    run use the same rounded weights and features)."""
    if precision not in _models:
        m = CaptionGenerator(echr_opt(), precision=precision)
        m.load_state_dict(helpers.state_dict("fp32" if precision == "fp32" else "bf16w"), strict=True)
        _models[precision] = m.to(DEV)
    return _models[precision]


def sd_of(precision):
    return helpers.state_dict("fp32" if precision == "fp32" else "bf16w")


def tf_check(eng, want_tok, per_video, precision, rows=None):
    """Teacher-forced log-probs of the staged batch along the ORACLE's tokens vs the oracle's, over the steps the
    reference evaluates (OldModel_NEW.py:98-130).  rows: only the first `rows` rows are compared."""
    tf = eng.teacher_forced(helpers.labels_of(want_tok).to(DEV))
    worst, r0 = 0.0, 0
    for n, c, want in per_video:
        if rows is not None and r0 >= rows:
            break
        if want is not None:
            got = tf[r0:r0 + n, :want.shape[1]].cpu()
            worst = max(worst, float((got - want).abs().max()))
        r0 += n
    assert worst < LOGP_HELD[precision] <= LOGP_ALLOWED[precision], worst
    return worst


# ------------------------------------------------------------------------------------------------ cfg2, benched schedule
def test_cfg2_f16_four_lanes_matches_oracle():
    precision = "f16"
    sd = sd_of(precision)
    vids, _ = synth.make_workload("cfg2", bf16_features=True)
    assert len(vids) == 64
    want_tok, want_lp, per_video = helpers.oracle_greedy_rows(sd, vids)
    m = model(precision)
    lay, d_c3d, d_tap, d_lda = helpers.batch_of(vids, DEV)
    _, h_c3d, h_tap, h_lda = helpers.batch_of(vids, pin=True)
    pipe = CaptionPipeline(m.engine, depth=4)
    try:
        # every lane sees the shape three times: eager, captured, replayed (bench.py's steady state)
        tickets = []
        for i in range(12):
            tickets.append(pipe.submit(lay, d_c3d, d_tap, d_lda, beam=1))
            if i >= 8:
                tok, lp, _ = tickets[-1].wait()
                torch.cuda.synchronize()
                exc, worst = helpers.check_greedy_rows(tok.cpu().numpy(), lp.cpu().numpy(), want_tok, want_lp, per_video,
                                                       LOGP_HELD[precision], "cfg2/f16/lane%d" % (i % 4))
                print("cfg2 f16 4 lanes (device buffers): %d of %d rows excused by a near tie, max |dlogp| %.2e"
                      % (exc, lay.n_events, worst))
        dev_tok = tok.cpu().clone()
        for i in range(8):
            t = pipe.submit(lay, h_c3d, h_tap, h_lda, beam=1)
        tok, lp, _ = t.wait()
        assert torch.equal(tok, dev_tok), "host-buffer path differs from the device-buffer path"
        # teacher-forced log-probs with the lanes' (shared-GPU) kernel schedule
        lane = pipe.lanes[1]
        lane.set_inputs(lay, d_c3d, d_tap, d_lda)
        lane.encode()
        worst = tf_check(lane, want_tok, per_video, precision)
        print("cfg2 f16 teacher-forced max |dlogp| vs oracle = %.2e (allowed %.0e)" % (worst, LOGP_ALLOWED[precision]))
    finally:
        pipe.close()


# ------------------------------------------------------------------------------------------------ >= 1024 rows
@pytest.mark.parametrize("precision", ["f16", "bf16w"])
def test_throughput_regime_matches_oracle(precision):
    """300 GT-style videos (~1190 rows >= 1024): persistent GEMM kernels + un-hoisted context k-blocks + streaming
    attention.  Oracle greedy + teacher-forced on the first 40 videos (rows are independent of the batch)."""
    sd = sd_of(precision)
    vids, _ = synth.make_workload("cfg2", n_videos=300, first_seed=5000, bf16_features=True)
    sub = vids[:40]
    want_tok, want_lp, per_video = helpers.oracle_greedy_rows(sd, sub)
    n_sub = want_tok.shape[0]
    m = model(precision)
    lay, c3d, tap, lda = helpers.batch_of(vids, DEV)
    assert lay.n_events >= 1024
    for _ in range(3):  # eager, captured, replayed
        tok, lp, _ = m.engine.caption_device(lay, c3d, tap, lda, beam=1)
    exc, worst = helpers.check_greedy_rows(tok[:n_sub].cpu().numpy(), lp[:n_sub].cpu().numpy(), want_tok, want_lp,
                                           per_video, LOGP_HELD[precision], "rows%d/%s" % (lay.n_events, precision))
    # teacher-forced: oracle tokens on the sub-sample rows, the path's own tokens elsewhere
    labels = tok.cpu().long()
    labels[:n_sub] = want_tok
    m.engine.set_inputs(lay, c3d, tap, lda)
    m.engine.encode()
    w = tf_check(m.engine, labels, per_video, precision, rows=n_sub)
    print("%d rows %s: %d of %d sampled rows excused, max |dlogp| %.2e, teacher-forced %.2e"
          % (lay.n_events, precision, exc, n_sub, worst, w))
    # the same videos alone (latency-regime kernels) give the same tokens except at near ties
    tok_small, _, _ = m.engine.caption_device(*helpers.batch_of(sub, DEV), beam=1)
    helpers.check_greedy_rows(tok_small.cpu().numpy(), want_lp.numpy(), want_tok, want_lp, per_video, 1.0, "sub-sample alone")


# ------------------------------------------------------------------------------------------------ cfg3
def _beam_case(precision, vids_gpu, vids_oracle, beam=3):
    sd = sd_of(precision)
    m = model(precision)
    lay, c3d, tap, lda = helpers.batch_of(vids_gpu, DEV)
    for _ in range(2):
        tok, lp, score = m.engine.caption_device(lay, c3d, tap, lda, beam=beam)
    parent, ptok = m.engine.beam_trace(beam)
    tok, lp, score, parent, ptok = tok.cpu(), lp.cpu(), score.cpu(), parent.cpu(), ptok.cpu()
    r0, stats = 0, [0, 0.0, 0, 0]
    for v in vids_oracle:
        n = v.n_events
        clean, worst, same = helpers.check_beam_replay(sd, v, beam, tok[r0:r0 + n].numpy(), lp[r0:r0 + n].numpy(),
                                                       score[r0:r0 + n].numpy(), parent[:, r0:r0 + n].numpy(),
                                                       ptok[:, r0:r0 + n].numpy(), LOGP_HELD[precision],
                                                       "beam%d/%s" % (beam, precision))
        stats = [stats[0] + clean, max(stats[1], worst), stats[2] + same, stats[3] + n]
        r0 += n
    return stats


@pytest.mark.parametrize("precision", ["fp32", "f16"])
def test_cfg3_video_relation_encoder_and_beam3(precision):
    """One cfg3 video: 100 SST-style proposals (100 x 100 pairs x 16 heads), greedy and beam = 3."""
    sd = sd_of(precision)
    v = synth.make_video(7000, T=100, n_events=100, style="sst", bf16_features=(precision != "fp32"))
    m = model(precision)
    want_tok, want_lp, per_video = helpers.oracle_greedy_rows(sd, [v])
    lay, c3d, tap, lda = helpers.batch_of([v], DEV)
    tok, lp, _ = m.engine.caption_device(lay, c3d, tap, lda, beam=1)
    ev = m.engine.event_context().cpu()
    ctx = O.build_context(sd, v.tap_feats, v.c3d_feats, v.lda_feats, v.ind_select_list, v.soi_select_list)
    ev_err = float((ev - ctx[1]).abs().max())
    assert ev_err < (2e-5 if precision == "fp32" else 2e-3), ev_err
    exc, worst = helpers.check_greedy_rows(tok.cpu().numpy(), lp.cpu().numpy(), want_tok, want_lp, per_video,
                                           LOGP_HELD[precision], "cfg3 video/" + precision)
    m.engine.set_inputs(lay, c3d, tap, lda)
    m.engine.encode()
    w = tf_check(m.engine, want_tok, per_video, precision)
    clean, regret, same, n = _beam_case(precision, [v], [v])
    print("cfg3 video %s: event ctx %.2e; greedy %d excused, |dlogp| %.2e, teacher-forced %.2e; beam 3: %d of %d events with "
          "zero-regret decisions throughout, max regret %.2e (near-tie bound 1e-3), %d identical to the spec's own search"
          % (precision, ev_err, exc, worst, w, clean, n, regret, same))


def test_cfg3_whole_batch_beam3_9600_rows():
    """BASELINE configs[2] whole: 32 videos x 100 proposals x beam 3 = 9600 decode rows (f16, the benched precision);
    oracle beam search on the first two videos."""
    vids, w = synth.make_workload("cfg3", bf16_features=True)
    assert len(vids) == 32 and w["beam"] == 3
    clean, regret, same, n = _beam_case("f16", vids, vids[:2])
    print("cfg3 whole (9600 rows) f16 beam 3: %d of %d sampled events with zero-regret decisions, max regret %.2e, %d identical "
          "to the spec's own search" % (clean, n, regret, same))


def test_beam3_fp32_in_the_throughput_regime():
    """Beam 3 over 4 cfg3 videos (1200 decode rows) in the two-plane precision: the persistent gate GEMM with the
    register-pipelined epilogue reading the cell state through the parent rows finalize_beam wrote; every decision
    of the first video replayed with the oracle's arithmetic."""
    vids, _ = synth.make_workload("cfg3", n_videos=4)
    clean, regret, same, n = _beam_case("fp32", vids, vids[:1])
    print("beam 3 fp32, 1200 rows: %d of %d events with zero-regret decisions, max regret %.2e, %d identical to the spec's own search"
          % (clean, n, regret, same))


# ------------------------------------------------------------------------------------------------ cfg4
@pytest.mark.parametrize("precision", ["fp32", "f16"])
def test_cfg4_video_long_spans(precision):
    """One cfg4 video: T = 400 clips, 200 SST-style proposals with spans up to K = 256 clips."""
    sd = sd_of(precision)
    v = synth.make_video(7100, T=400, n_events=200, style="sst", bf16_features=(precision != "fp32"))
    assert max(e - s for s, e in v.soi_select_list) > 200
    m = model(precision)
    want_tok, want_lp, per_video = helpers.oracle_greedy_rows(sd, [v])
    lay, c3d, tap, lda = helpers.batch_of([v], DEV)
    tok, lp, _ = m.engine.caption_device(lay, c3d, tap, lda, beam=1)
    ev = m.engine.event_context().cpu()
    ctx = O.build_context(sd, v.tap_feats, v.c3d_feats, v.lda_feats, v.ind_select_list, v.soi_select_list)
    ev_err = float((ev - ctx[1]).abs().max())
    assert ev_err < (2e-5 if precision == "fp32" else 2e-3), ev_err
    exc, worst = helpers.check_greedy_rows(tok.cpu().numpy(), lp.cpu().numpy(), want_tok, want_lp, per_video,
                                           LOGP_HELD[precision], "cfg4 video/" + precision)
    m.engine.set_inputs(lay, c3d, tap, lda)
    m.engine.encode()
    w = tf_check(m.engine, want_tok, per_video, precision)
    print("cfg4 video %s: event ctx %.2e; greedy %d of 200 excused, |dlogp| %.2e, teacher-forced %.2e"
          % (precision, ev_err, exc, worst, w))


def test_cfg4_whole_batch_matches_single_videos():
    """BASELINE configs[3] whole (16 videos, 3200 rows, f16) against the same videos decoded one per call (whose
    kernels the test above pins to the oracle): tokens equal except at near ties of the oracle on video 0."""
    vids, _ = synth.make_workload("cfg4", bf16_features=True)
    m = model("f16")
    lay, c3d, tap, lda = helpers.batch_of(vids, DEV)
    for _ in range(2):
        tok, lp, _ = m.engine.caption_device(lay, c3d, tap, lda, beam=1)
    tok, lp = tok.cpu(), lp.cpu()
    want_tok, want_lp, per_video = helpers.oracle_greedy_rows(sd_of("f16"), vids[:1])
    exc, worst = helpers.check_greedy_rows(tok[:200].numpy(), lp[:200].numpy(), want_tok, want_lp, per_video,
                                           LOGP_HELD["f16"], "cfg4 whole")
    print("cfg4 whole f16: video 0 vs oracle: %d of 200 excused, |dlogp| %.2e" % (exc, worst))
    single = m.engine.caption_device(*helpers.batch_of(vids[3:4], DEV), beam=1)[0].cpu()
    agree = float((single == tok[600:800]).all(1).float().mean())
    assert agree > 0.9, agree


def test_throughput_regime_with_a_vocabulary_below_one_tile():
    """>= 1024 rows with a vocabulary narrower than the persistent kernel's 256-column tile: the logit GEMM stays on the
    one-tile-per-CTA kernel, so the next step's query GEMM cannot ride in it and the decode must fall back to the
    schedule with the query GEMM as its own launch.  Checked against the same videos decoded in chunks below 1024 rows
    (the latency-regime kernels, pinned to the oracle by test_engine_gpu.py)."""
    opt = echr_opt(CG_seq_length=6, CG_vocab_size=200)
    sd = synth.round_state_dict_bf16(synth.init_state_dict(opt, seed=11))
    m = CaptionGenerator(opt, precision="f16")
    m.load_state_dict(sd, strict=True)
    m = m.to(DEV)
    vids, _ = synth.make_workload("cfg5", n_videos=280, bf16_features=True)
    lay, c3d, tap, lda = helpers.batch_of(vids, DEV)
    assert lay.n_events >= 1024
    for _ in range(2):  # eager on first sight, the captured graph on the second
        tok, lp, _ = m.engine.caption_device(lay, c3d, tap, lda, beam=1)
    tok, lp = tok.cpu(), lp.cpu()
    assert tuple(tok.shape) == (lay.n_events, 6) and int(tok.max()) <= 200
    row, same, worst = 0, 0, 0.0
    for a in range(0, 280, 70):
        part = vids[a:a + 70]
        t2, l2, _ = m.engine.caption_device(*helpers.batch_of(part, DEV), beam=1)
        t2, l2 = t2.cpu(), l2.cpu()
        n = t2.shape[0]
        eq = (t2 == tok[row:row + n]).all(1)
        same += int(eq.sum())
        if eq.any():
            worst = max(worst, float((l2[eq] - lp[row:row + n][eq]).abs().max()))
        row += n
    assert row == lay.n_events
    assert same >= 0.97 * row, (same, row)      # the rest: near ties resolved differently by the two GEMM kernels
    assert worst < LOGP_HELD["f16"], worst


# ------------------------------------------------------------------------------------------------ topN = 1000
def test_one_video_with_1000_proposals():
    """The reference's default selection hands up to topN = 1000 proposals of one video to the caption call
    (eval_utils.py:21,107-115): relation encoder over 1000 x 1000 pairs (key chunks + query blocks of the relation
    attention, 7813 tiles of the fused pair kernel) against the oracle evaluated in query tiles, then the decode of
    the first 24 proposals against the oracle's decode of those rows."""
    precision = "fp32"
    sd = sd_of(precision)
    v = synth.make_video(7200, T=400, n_events=1000, style="sst")
    m = model(precision)
    lay, c3d, tap, lda = helpers.batch_of([v], DEV)
    tok, lp, _ = m.engine.caption_device(lay, c3d, tap, lda, beam=1)
    ev = m.engine.event_context().cpu()
    soi, ind = v.soi_select_list, v.ind_select_list
    pooled = torch.stack([v.c3d_feats[s:e].mean(0) for s, e in soi], 0)
    feats = torch.cat([pooled, v.tap_feats[torch.as_tensor(ind)]], 1)
    want_ev = O.relation_encode_tiled(sd, feats, soi, tile=50)
    err = float((ev - want_ev).abs().max())
    assert err < 2e-5, err
    n_sub = 24
    lmax = max(e - s for s, e in soi[:n_sub])
    clip = torch.zeros(n_sub, lmax, 500)
    mask = torch.zeros(n_sub, lmax)
    for i, (s, e) in enumerate(soi[:n_sub]):
        clip[i, :e - s] = v.c3d_feats[s:e]
        mask[i, :e - s] = 1
    s_ref, l_ref = O.greedy_decode(sd, v.lda_feats, want_ev[:n_sub], clip, mask)
    tf = O.teacher_forced_logprobs(sd, v.lda_feats, want_ev[:n_sub], clip, mask, helpers.labels_of(s_ref))
    c = s_ref.shape[1]
    want_tok = torch.zeros(n_sub, 30, dtype=torch.long)
    want_lp = torch.zeros(n_sub, 30)
    want_tok[:, :c], want_lp[:, :c] = s_ref, l_ref
    exc, worst = helpers.check_greedy_rows(tok[:n_sub].cpu().numpy(), lp[:n_sub].cpu().numpy(), want_tok, want_lp,
                                           [(n_sub, c, tf)], LOGP_HELD[precision], "topN=1000")
    print("1000 proposals: event ctx max |d| %.2e; first %d rows: %d excused, |dlogp| %.2e" % (err, n_sub, exc, worst))
