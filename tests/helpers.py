"""Shared helpers of the parity tests."""
import os

import numpy as np
import torch

from echr_b200 import synth
from echr_b200.config import echr_opt

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# name -> (video seed, T, n_events, style); must match tests/golden/make_golden.py
CASES = {
    "cfg1": (0, 100, 3, "gt"),
    "gt5": (1, 100, None, "gt"),
    "sst12": (2, 100, 12, "sst"),
    "long30": (3, 400, 30, "sst"),
}
NEAR_TIE = 1e-3  # BASELINE.json north_star: token mismatches are excused where the top-2 gap is below this

_sd_cache = {}


def state_dict(kind="fp32"):
    if kind not in _sd_cache:
        sd = synth.init_state_dict(echr_opt(), seed=0)
        if kind == "bf16w":
            sd = synth.round_state_dict_bf16(sd)
        _sd_cache[kind] = sd
    return _sd_cache[kind]


def case_video(name, kind="fp32"):
    seed, T, n, style = CASES[name]
    return synth.make_video(seed, T=T, n_events=n, style=style, bf16_features=(kind == "bf16w"))


def golden(name, kind="fp32"):
    return np.load(os.path.join(GOLDEN_DIR, "%s_%s.npz" % (name, kind)))


def check_tokens(got, want, top2, what=""):
    """got/want int [N,S]; top2 [N,S,2] = the oracle's two best log-probs at every step along `want`.
    Rows must agree up to the first step where the oracle's top-2 gap is below NEAR_TIE; returns the
    number of rows excused that way."""
    got, want = np.asarray(got), np.asarray(want)
    assert got.shape == want.shape, (got.shape, want.shape)
    excused = 0
    for r in range(want.shape[0]):
        diff = np.nonzero(got[r] != want[r])[0]
        if diff.size == 0:
            continue
        t = int(diff[0])
        gap = float(top2[r, t, 0] - top2[r, t, 1])
        assert gap < NEAR_TIE, "%s row %d diverges at step %d (got %d want %d) with top-2 gap %.3e" % (
            what, r, t, got[r, t], want[r, t], gap)
        excused += 1
    return excused


def video_tuple(v, device=None):
    mv = (lambda t: t.to(device)) if device is not None else (lambda t: t)
    return (mv(v.tap_feats), mv(v.c3d_feats), mv(v.lda_feats), list(v.ind_select_list),
            [list(s) for s in v.soi_select_list])


def labels_of(seq):
    """[N,S] tokens -> [N,S+1] labels with the BOS column the teacher-forced forward expects."""
    seq = torch.as_tensor(np.asarray(seq)).long()
    return torch.cat([torch.zeros(seq.shape[0], 1, dtype=torch.long), seq], 1)


def batch_of(vids, device=None, pin=False):
    """(BatchLayout, c3d, tap, lda) of a list of synth videos, concatenated."""
    from echr_b200.engine import BatchLayout
    lay = BatchLayout([v.c3d_feats.shape[0] for v in vids], [v.ind_select_list for v in vids],
                      [v.soi_select_list for v in vids])
    ts = (torch.cat([v.c3d_feats for v in vids]).contiguous(), torch.cat([v.tap_feats for v in vids]).contiguous(),
          torch.stack([v.lda_feats for v in vids]).contiguous())
    if pin:
        ts = tuple(t.pin_memory() for t in ts)
    if device is not None:
        ts = tuple(t.to(device) for t in ts)
    return (lay,) + ts


def oracle_greedy_rows(sd, vids, seq_length=30):
    """Oracle greedy decode + teacher-forced log-probs of a list of videos (one video per call, like the
    reference), padded to full width: (tokens int64 [rows,S], logp [rows,S], per_video) with per_video a list of
    (events, columns the reference returned, teacher-forced log-probs [events, columns, V+1] or None)."""
    from oracle import echr_oracle as O
    toks, lps, per_video = [], [], []
    for v in vids:
        ctx = O.build_context(sd, v.tap_feats, v.c3d_feats, v.lda_feats, v.ind_select_list, v.soi_select_list)
        s, l = O.greedy_decode(sd, *ctx, seq_length=seq_length)
        n = v.n_events
        t = torch.zeros(n, seq_length, dtype=torch.long)
        p = torch.zeros(n, seq_length)
        c, tf = 0, None
        if len(s):
            c = s.shape[1]
            t[:, :c], p[:, :c] = s, l
            tf = O.teacher_forced_logprobs(sd, *ctx, labels_of(s))
        toks.append(t)
        lps.append(p)
        per_video.append((n, c, tf))
    return torch.cat(toks), torch.cat(lps), per_video


def check_greedy_rows(got_tok, got_lp, want_tok, want_lp, per_video, lp_tol, what=""):
    """Rows of the CUDA path (full width, tokens masked after EOS) against oracle_greedy_rows(): tokens under the
    near-tie rule over the columns the reference returned; returned log-probs of rows that did not diverge
    within lp_tol.  Returns (excused rows, max |dlogp| over the compared entries)."""
    got_tok, got_lp = np.asarray(got_tok), np.asarray(got_lp)
    want_tok, want_lp = want_tok.numpy(), want_lp.numpy()
    excused, worst, r0 = 0, 0.0, 0
    for n, c, tf in per_video:
        if tf is None:  # the reference returned ([], []): every row's first token is EOS
            assert (got_tok[r0:r0 + n, 0] == 0).all(), what
            r0 += n
            continue
        top2 = tf.topk(2, dim=2)[0].numpy()
        for r in range(n):
            g, w = got_tok[r0 + r, :c], want_tok[r0 + r, :c]
            diff = np.nonzero(g != w)[0]
            if diff.size:
                t = int(diff[0])
                gap = float(top2[r, t, 0] - top2[r, t, 1])
                assert gap < NEAR_TIE, "%s row %d diverges at step %d (got %d want %d) with top-2 gap %.3e" % (
                    what, r0 + r, t, g[t], w[t], gap)
                excused += 1
                continue
            d = float(np.abs(got_lp[r0 + r, :c] - want_lp[r0 + r, :c]).max())
            worst = max(worst, d)
            assert d < lp_tol, "%s row %d: |dlogp| %.3e" % (what, r0 + r, d)
        r0 += n
    return excused, worst


def beam_len(row):
    """Tokens of a zero-padded beam hypothesis including its EOS."""
    z = np.nonzero(np.asarray(row) == 0)[0]
    return int(z[0]) + 1 if z.size else len(row)


def check_beam_replay(sd, v, beam, got_tok, got_lp, got_score, parent, tok, lp_tol, what=""):
    """Beam rule (BASELINE north_star, applied to the search's decisions): the CUDA path's per-step decisions
    (echr_get_beam_trace) are replayed with the oracle's arithmetic (oracle.replay_beam_trace).  Every step must
    keep a best-`beam` set of proposals up to a near tie (the worst kept proposal within NEAR_TIE of the beam-th
    best one), from live parents, without duplicates; the returned hypothesis must be the best finished one along
    the trace (or within NEAR_TIE of it); its log-probs / score must be the oracle's within lp_tol.
    Returns (rows whose decisions all had zero regret, max regret, rows equal to the spec's own search)."""
    from oracle import echr_oracle as O
    ctx = O.build_context(sd, v.tap_feats, v.c3d_feats, v.lda_feats, v.ind_select_list, v.soi_select_list)
    r = O.replay_beam_trace(sd, *ctx, parent, tok, beam)
    assert not r["invalid"], "%s: invalid decisions (dead parent / duplicate proposal) at %s" % (what, r["invalid"][:5])
    regret = r["regret"].numpy()  # [S, N]
    worst = float(regret.max())
    assert worst < NEAR_TIE, "%s: a step kept a proposal %.3e below the beam-th best one (step %d, event %d)" % (
        what, worst, *np.unravel_index(regret.argmax(), regret.shape))
    got_tok, got_lp, got_score = np.asarray(got_tok), np.asarray(got_lp), np.asarray(got_score)
    for e in range(got_tok.shape[0]):
        match = [f for f in r["finished"][e] if (f[1].numpy() == got_tok[e]).all()]
        assert match, "%s event %d: the returned hypothesis is not a finished hypothesis of its own trace" % (what, e)
        sc, _, lps = match[0]
        assert float(r["best_score"][e]) - sc < NEAR_TIE, "%s event %d: returned hypothesis loses to the best finished one by %.3e" % (
            what, e, float(r["best_score"][e]) - sc)
        L = beam_len(got_tok[e])
        assert np.abs(lps.numpy()[:L] - got_lp[e, :L]).max() < lp_tol, (what, e)
        assert abs(sc - float(got_score[e])) < 30 * lp_tol, (what, e, sc, float(got_score[e]))
    t_ref = O.beam_search(sd, *ctx, beam=beam)[0].numpy()
    return int((regret.max(0) == 0).sum()), worst, int((t_ref == got_tok).all(1).sum())
