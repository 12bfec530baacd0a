"""Proposal selection.  CPU part: the oracle restatement against the outputs of the
reference's own gettop1000 / gettop1000_nms / gettopN_nms (tests/golden/proposals.npz).  GPU part: the CUDA
kernels through the C-ABI and through the reference-shaped host functions, bit-exact against both."""
import os

import numpy as np
import pytest
import torch

import proposals_cases as C
from oracle import proposals_oracle as PO

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "proposals.npz"))


# ----------------------------------------------------------------------------- oracle vs the reference (CPU)
@pytest.mark.parametrize("c", range(len(C.TOPN_CASES)))
def test_oracle_topn_matches_reference(c):
    s, m, gts, topn, thres = C.topn_inputs(c)
    ind, start, end, conf = PO.topn_select(s, m, topn=topn, val_score_thres=thres)
    pre = "topn%d_" % c
    assert np.array_equal(ind, G[pre + "ind"])
    assert np.array_equal(np.stack([start, end], 1), G[pre + "feat"])
    assert np.array_equal(conf.astype(np.float64), G[pre + "conf"])
    assert np.array_equal(gts[ind, ind - start], G[pre + "cg"])


@pytest.mark.parametrize("c", range(len(C.NMS_CASES)))
def test_oracle_nms_matches_reference(c):
    s, gts, ov, topn = C.nms_inputs(c)
    pick, props, sc = PO.nms_select(s, overlap=ov, topn=topn)
    pre = "nms%d_" % c
    assert np.array_equal(props, G[pre + "props"])
    assert np.array_equal(props[:, 1] - 1, G[pre + "ind"])
    assert np.array_equal(sc, G[pre + "sc"])
    n, k = PO.enumerate_proposals(*s.shape)
    assert np.array_equal(gts[n[pick], k[pick]], G[pre + "pg"])


@pytest.mark.parametrize("c", range(len(C.RERANK_CASES)))
def test_oracle_rerank_matches_reference(c):
    props, ps, ss, ov, topn = C.rerank_inputs(c)
    assert PO.rerank_nms(props, ps, ss, nms_overlap=ov, topn=topn) == G["rr%d_pick" % c].tolist()


def test_proposal_enumeration_count_matches_library_formula():
    for T, K in ((0, 4), (1, 4), (2, 4), (5, 4), (6, 4), (100, 256), (257, 256), (400, 256)):
        n, k = PO.enumerate_proposals(T, K)
        assert n.size == sum(min(i, K) for i in range(T))
        assert (k < np.minimum(n, K)).all()


# ----------------------------------------------------------------------------- CUDA kernels (GPU)
@pytest.mark.gpu
@pytest.mark.parametrize("c", range(len(C.TOPN_CASES)))
def test_gpu_gettop1000_matches_reference(c):
    from echr_b200 import proposals as P
    s, m, gts, topn, thres = C.topn_inputs(c)
    ind, feat, cg, ts, conf = P.gettop1000(torch.from_numpy(s).cuda(), m, gts, 120.0, C.featstamp_to_time,
                                           val_score_thres=thres, topN=topn)
    pre = "topn%d_" % c
    assert ind == G[pre + "ind"].tolist()
    assert feat == G[pre + "feat"].tolist()
    assert [int(x) for x in cg] == G[pre + "cg"].tolist()
    assert conf == G[pre + "conf"].tolist()  # float32 values widened to python floats, bit-exact
    assert np.array_equal(np.asarray(ts, dtype=np.float64).reshape(-1, 2), G[pre + "ts"])


@pytest.mark.gpu
@pytest.mark.parametrize("c", range(len(C.NMS_CASES)))
def test_gpu_gettop1000_nms_matches_reference(c):
    from echr_b200 import proposals as P
    s, gts, ov, topn = C.nms_inputs(c)
    ind, props, pg, ts, sc = P.gettop1000_nms(torch.from_numpy(s).cuda(), None, gts, 90.0, C.featstamp_to_time,
                                              overlap=ov, topN=topn)
    pre = "nms%d_" % c
    assert np.array_equal(ind, G[pre + "ind"]) and np.array_equal(props, G[pre + "props"])
    assert np.array_equal(pg, G[pre + "pg"]) and np.array_equal(sc, G[pre + "sc"])
    assert np.array_equal(np.asarray(ts, dtype=np.float64).reshape(-1, 2), G[pre + "ts"])


@pytest.mark.gpu
@pytest.mark.parametrize("c", range(len(C.RERANK_CASES)))
def test_gpu_gettopN_nms_matches_reference(c):
    from echr_b200 import proposals as P
    props, ps, ss, ov, topn = C.rerank_inputs(c)
    nms_props, nms_scores, pick = P.gettopN_nms(props.tolist(), ps.tolist(), ss.tolist(), nms_overlap=ov, topN=topn)
    want = G["rr%d_pick" % c]
    assert [int(i) for i in pick] == want.tolist()
    assert np.array_equal(nms_props, props[want]) and np.array_equal(nms_scores, ps[want])


@pytest.mark.gpu
def test_gpu_batched_selection_matches_oracle_per_video():
    """Many ragged videos per launch (one CTA each), device tensors out: equal to the oracle video by video;
    includes a 1-clip and an empty video, masked-out matrices (threshold 0 admits zero cells) and a video with
    more than 65k proposals."""
    from echr_b200 import proposals as P
    rng = np.random.RandomState(5)
    K = 64
    Ts = [37, 1, 0, 100, 64, 65, 66, 1200, 5]
    off = np.concatenate([[0], np.cumsum(Ts)])
    s = np.concatenate([C.distinct_scores(rng, T, K) if T else np.zeros((0, K), np.float32) for T in Ts])
    m = (rng.rand(*s.shape) < 0.3).astype(np.float32)
    d_s = torch.from_numpy(s).cuda()
    ind, start, end, conf, count = [t.cpu().numpy() for t in P.select_topn(d_s, torch.from_numpy(m).cuda(), off, topN=300)]
    for v, T in enumerate(Ts):
        if T == 0:
            assert count[v] == 0
            continue
        wi, ws, we, wc = PO.topn_select(s[off[v]:off[v + 1]], m[off[v]:off[v + 1]], topn=300)
        c = count[v]
        assert c == wi.size
        assert np.array_equal(ind[v, :c], wi) and np.array_equal(start[v, :c], ws) and np.array_equal(end[v, :c], we)
        assert np.array_equal(conf[v, :c], wc)
    ind, start, end, sc, pick, count = [t.cpu().numpy() for t in P.select_nms(d_s, off, overlap=0.6, topN=150)]
    for v, T in enumerate(Ts):
        if T < 2:
            assert count[v] == 0
            continue
        wp, wprops, wsc = PO.nms_select(s[off[v]:off[v + 1]], overlap=0.6, topn=150)
        c = count[v]
        assert c == wp.size
        assert np.array_equal(pick[v, :c], wp) and np.array_equal(start[v, :c], wprops[:, 0])
        assert np.array_equal(end[v, :c], wprops[:, 1]) and np.array_equal(ind[v, :c], wprops[:, 1] - 1)
        assert np.array_equal(sc[v, :c].astype(np.float64), wsc)


@pytest.mark.gpu
def test_gpu_selection_feeds_the_caption_path():
    """SST-style flow without the host loop: scores [T,K] on the device -> select_topn -> the (ind, start, end)
    lists the caption path takes (CaptionGenerator.py:121, eval_utils.py:114-129)."""
    from echr_b200 import proposals as P
    import helpers
    from echr_b200 import synth
    from echr_b200.caption_generator import CaptionGenerator
    from echr_b200.config import echr_opt
    rng = np.random.RandomState(9)
    T, K = 60, 32
    s = C.distinct_scores(rng, T, K)
    m = (np.arange(T)[:, None] >= np.arange(K)[None, :]).astype(np.float32)
    ind, feat, _, _, conf = P.gettop1000(torch.from_numpy(s).cuda(), m, [], 60.0, C.featstamp_to_time, topN=12)
    wi, ws, we, _ = PO.topn_select(s, m, topn=12)
    assert ind == wi.tolist() and feat == np.stack([ws, we], 1).tolist() and len(ind) == 12
    v = synth.make_video(3, T=T, n_events=3)
    model = CaptionGenerator(echr_opt(), precision="fp32")
    model.load_state_dict(helpers.state_dict("fp32"))
    model = model.cuda()
    seq, logp = model(v.tap_feats.cuda(), v.c3d_feats.cuda(), v.lda_feats.cuda(), [], ind, feat, mode="eval")
    assert seq.shape[0] == 12 and torch.isfinite(logp).all()
