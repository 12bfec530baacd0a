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


# ----------------------------------------------------------------------------- size-independent properties (CPU)
@pytest.mark.parametrize("seed", range(6))
def test_oracle_selection_properties(seed):
    """Invariants the domain offers, on shapes the golden cases do not cover."""
    rng = np.random.RandomState(50 + seed)
    T, K = int(rng.randint(2, 90)), int(rng.choice([4, 16, 64, 256]))
    s = C.distinct_scores(rng, T, K)
    m = (np.arange(T)[:, None] >= np.arange(K)[None, :]).astype(np.float32)
    topn = int(rng.randint(1, 300))
    ind, start, end, conf = PO.topn_select(s, m, topn=topn)
    # every emitted cell is a valid (n >= k) cell at or above the topn-th largest masked score, in row-major order
    p = s * m
    thr = np.sort(p.reshape(-1))[-min(p.size, topn)]
    assert (conf >= thr).all() and (ind - start >= 0).all() and (end == ind + 1).all()
    order = ind * K + (ind - start)
    assert (np.diff(order) > 0).all()
    assert len(ind) == int(((p >= max(thr, 0)) & (m > 0)).sum() + ((p >= max(thr, 0)) & (m == 0) & (np.arange(T)[:, None] >= np.arange(K)[None, :])).sum())
    # NMS: scores come out in descending order, no two picks overlap by more than the threshold, and every
    # proposal left out overlaps an earlier-picked, better-scored one (unless the pick budget ran out)
    ov = float(rng.choice([0.3, 0.6, 0.8]))
    budget = int(rng.randint(1, 200))
    pick, props, sc = PO.nms_select(s, overlap=ov, topn=budget)
    assert (np.diff(sc) < 0).all()

    def iou(a, b):
        wh = max(0.0, min(a[1], b[1]) - max(a[0], b[0]) + 1.0)
        return wh / ((a[1] - a[0] + 1.0) + (b[1] - b[0] + 1.0) - wh)
    for i in range(len(props)):
        for j in range(i):
            assert iou(props[i], props[j]) <= ov
    if len(pick) < budget:
        n, k = PO.enumerate_proposals(T, K)
        picked = set(pick.tolist())
        for q in rng.choice(n.size, size=min(40, n.size), replace=False):
            if int(q) in picked:
                continue
            box = (n[q] - k[q], n[q] + 1)
            assert any(iou(box, props[j]) > ov and sc[j] > s[n[q], k[q]] for j in range(len(props)))
    # idempotence: selecting again among the picks alone changes nothing
    keep = np.full_like(s, -1.0)
    n, k = PO.enumerate_proposals(T, K)
    keep[n[pick], k[pick]] = s[n[pick], k[pick]]
    pick2, props2, _ = PO.nms_select(keep, overlap=ov, topn=len(pick))
    assert np.array_equal(props2, props)


@pytest.mark.gpu
def test_gpu_selection_properties_at_full_size():
    """BASELINE-size matrices (64 videos x T=400 x K=256, 69.5k proposals per video) are too slow for the numpy
    oracle: check what must hold at any size — descending scores, pairwise overlap of the picks <= threshold,
    the pick budget, top-N threshold / order / count — and the oracle itself on two of the videos."""
    from echr_b200 import proposals as P
    rng = np.random.RandomState(8)
    V, T, K, ov, topn = 64, 400, 256, 0.7, 300
    s = np.stack([C.distinct_scores(rng, T, K) for _ in range(V)]).reshape(V * T, K)
    off = np.arange(V + 1) * T
    d_s = torch.from_numpy(s).cuda()
    ind, start, end, sc, pick, count = P.select_nms(d_s, off, overlap=ov, topN=topn)
    assert int(count.min()) >= 1 and int(count.max()) <= topn
    valid = torch.arange(sc.shape[1], device="cuda")[None, :] < count[:, None]
    assert bool(((sc[:, 1:] < sc[:, :-1]) | ~valid[:, 1:]).all())
    a0, a1 = start.double(), end.double()
    wh = (torch.minimum(a1[:, :, None], a1[:, None, :]) - torch.maximum(a0[:, :, None], a0[:, None, :]) + 1.0).clamp(min=0)
    area = a1 - a0 + 1.0
    o = wh / (area[:, :, None] + area[:, None, :] - wh)
    pair = valid[:, :, None] & valid[:, None, :] & ~torch.eye(sc.shape[1], dtype=torch.bool, device="cuda")[None]
    assert float(o[pair].max()) <= ov
    assert bool((ind == end - 1)[valid].all())
    got_score = d_s.view(V, T, K)[torch.arange(V, device="cuda")[:, None], ind.long(), (ind - start).long()]
    assert torch.equal(got_score[valid], sc[valid])
    for v in (0, V - 1):
        wp, wprops, wsc = PO.nms_select(s[off[v]:off[v + 1]], overlap=ov, topn=topn)
        c = int(count[v])
        assert c == wp.size and np.array_equal(pick[v, :c].cpu().numpy(), wp)
    # top-N
    m = np.tile((np.arange(T)[:, None] >= np.arange(K)[None, :]).astype(np.float32), (V, 1))
    ti, ts, te, tc, tcount = P.select_topn(d_s, torch.from_numpy(m).cuda(), off, topN=1000)
    assert bool((tcount == 1000).all())  # distinct scores, mask = the valid cells: exactly topN cells qualify
    p = torch.from_numpy(s * m).cuda().view(V, -1)
    thr = p.topk(1000, dim=1)[0][:, -1]
    assert bool((tc[:, :1000] >= thr[:, None]).all())
    order = ti[:, :1000].long() * K + (ti - ts)[:, :1000].long()
    assert bool((order[:, 1:] > order[:, :-1]).all())
    assert torch.equal(p.gather(1, order), tc[:, :1000])
