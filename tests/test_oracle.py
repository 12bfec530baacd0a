"""CPU tests: the oracle restatement against the golden vectors produced by the reference itself,
and the invariants the reference pins for the (unpinned) beam spec."""
import numpy as np
import pytest
import torch

import helpers
from echr_b200 import synth
from oracle import echr_oracle as O

torch.set_grad_enabled(False)


@pytest.mark.parametrize("kind", ["fp32", "bf16w"])
@pytest.mark.parametrize("name", ["cfg1", "gt5", "sst12"])
def test_oracle_matches_reference_golden(name, kind):
    sd, v, g = helpers.state_dict(kind), helpers.case_video(name, kind), helpers.golden(name, kind)
    video, event, clip, mask = O.build_context(sd, v.tap_feats, v.c3d_feats, v.lda_feats, v.ind_select_list,
                                               v.soi_select_list)
    assert np.abs(event.numpy() - g["event"]).max() < 2e-6
    seq, logp = O.greedy_decode(sd, video, event, clip, mask)
    assert np.array_equal(seq.numpy(), g["seq"])
    assert np.abs(logp.numpy() - g["logp"]).max() < 2e-5
    labels = torch.cat([torch.zeros(seq.shape[0], 1, dtype=torch.long), seq], 1)
    tf = O.teacher_forced_logprobs(sd, video, event, clip, mask, labels)
    steps = list(g["tf_steps"])
    ne = g["tf_logprobs"].shape[0]
    assert np.abs(tf[:ne][:, steps].numpy() - g["tf_logprobs"]).max() < 2e-5
    assert np.abs(tf.topk(2, dim=2)[0].numpy() - g["tf_top2"]).max() < 2e-5


def test_position_embedding_layout():
    soi = [[0, 10], [5, 6], [20, 60]]
    pm = O.position_matrix(soi)
    pe = O.position_embedding(pm)
    assert pe.shape == (3, 3, 512)
    f = 100.0 / np.power(10000.0, np.arange(128) / 128.0)
    assert np.allclose(pe[0, 1, :128], np.sin(pm[0, 1, 0] * f))
    assert np.allclose(pe[0, 1, 128:256], np.cos(pm[0, 1, 0] * f))
    assert np.allclose(pe[0, 1, 256:384], np.sin(pm[0, 1, 1] * f))
    assert np.allclose(pe[0, 1, 384:], np.cos(pm[0, 1, 1] * f))
    assert pm[0, 0, 0] == 1e-3 and pm[0, 0, 1] == 0.0  # clamp on the diagonal


def test_beam_invariants():
    """What the reference does pin for beam search: beam=1 is its greedy decode, every returned
    hypothesis' per-step log-probs are the teacher-forced ones, and a wider beam never scores lower."""
    sd, v, g = helpers.state_dict("fp32"), helpers.case_video("cfg1"), helpers.golden("cfg1")
    ctx = O.build_context(sd, v.tap_feats, v.c3d_feats, v.lda_feats, v.ind_select_list, v.soi_select_list)
    t1, l1, s1 = O.beam_search(sd, *ctx, beam=1)
    assert np.array_equal(t1.numpy(), g["seq"])
    assert np.abs(l1.numpy() - g["logp"]).max() < 2e-5
    t3, l3, s3 = O.beam_search(sd, *ctx, beam=3)
    assert bool((s3 >= s1 - 1e-4).all())
    assert np.allclose(l3.sum(1).numpy(), s3.numpy(), atol=1e-3)
    labels = torch.cat([torch.zeros(t3.shape[0], 1, dtype=torch.long), t3], 1)
    tf = O.teacher_forced_logprobs(sd, *ctx, labels)
    picked = torch.gather(tf, 2, t3[:, : tf.shape[1]].unsqueeze(2)).squeeze(2)
    assert np.abs(picked.numpy() - l3[:, : tf.shape[1]].numpy()).max() < 1e-4


def test_greedy_empty_and_trim():
    """A row whose first argmax is token 0 is finished at once; when every row is, the reference
    returns ([], []) (OldModel_NEW.py:185-186).  Force it with a logit bias on token 0."""
    sd = dict(helpers.state_dict("fp32"))
    b = sd["lm_model.logit.bias"].clone()
    b[0] = 50.0
    sd["lm_model.logit.bias"] = b
    v = helpers.case_video("cfg1")
    seq, logp = O.caption_video(sd, v.tap_feats, v.c3d_feats, v.lda_feats, v.ind_select_list, v.soi_select_list)
    assert seq == [] and logp == []


def test_regime_helpers_accept_the_oracle_and_reject_a_wrong_row():
    """The checkers of tests/test_0_regimes_gpu.py on CPU: the oracle's own output passes; a row changed at a step
    with a clear top-2 gap / a clearly worse beam hypothesis is refused."""
    import pytest
    sd = helpers.state_dict("fp32")
    vids = [synth.make_video(300 + i, n_events=n) for i, n in enumerate((3, 4))]
    tok, lp, per_video = helpers.oracle_greedy_rows(sd, vids)
    exc, worst = helpers.check_greedy_rows(tok.numpy(), lp.numpy(), tok, lp, per_video, 1e-6)
    assert exc == 0 and worst == 0.0
    bad = tok.numpy().copy()
    top2 = per_video[0][2].topk(2, dim=2)[0].numpy()
    t = int(np.argmax(top2[0, :, 0] - top2[0, :, 1]))  # the clearest decision of row 0
    bad[0, t] = (bad[0, t] + 1) % 5748
    with pytest.raises(AssertionError):
        helpers.check_greedy_rows(bad, lp.numpy(), tok, lp, per_video, 1e-6)
    # beam: a trace of the spec's own decisions replays with zero regret; a trace that keeps a clearly worse proposal
    # is measured as such
    v = vids[1]
    ctx = O.build_context(sd, v.tap_feats, v.c3d_feats, v.lda_feats, v.ind_select_list, v.soi_select_list)
    t_ref, l_ref, s_ref, trace = O.beam_search(sd, *ctx, beam=3, return_trace=True)
    S, n, B = 30, 4, 3
    parent, tok = np.zeros((S, n, B), int), np.zeros((S, n, B), int)
    prev = None
    for t in range(S):
        kt = trace[t]["kept_tok"].numpy()
        tok[t] = kt[:, :, t]
        for e in range(n):
            for k in range(B):
                if t > 0:
                    parent[t, e, k] = [b for b in range(B) if (prev[e, b] == kt[e, k, :t]).all()][0]
        prev = kt
    r = O.replay_beam_trace(sd, *ctx, parent, tok, 3)
    assert float(r["regret"].max()) == 0.0 and not r["invalid"] and torch.equal(r["best_tok"], t_ref)
    clean, worst, same = helpers.check_beam_replay(sd, v, 3, t_ref.numpy(), l_ref.numpy(), s_ref.numpy(), parent, tok, 1e-5)
    assert (clean, worst, same) == (4, 0.0, 4)
    bad = tok.copy()
    bad[3, 0, 2] = 17  # an arbitrary token instead of the third-best proposal
    assert float(O.replay_beam_trace(sd, *ctx, parent, bad, 3)["regret"][3, 0]) > 1e-2
