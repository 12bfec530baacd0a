"""TEST INFRASTRUCTURE (CPU restatement, numpy): proposal selection between the proposal module and the
caption path.  Only tests/, smoke() and the CPU-baseline leg of the measurement tools may
import this; the product path is echr_b200/proposals.py -> CUDA.

Restates, with different control flow (mask arrays instead of shrinking index lists, a vectorised
compaction instead of the double loop):
  * topn_select   eval_utils.py:263-292  gettop1000
  * nms_select    eval_utils.py:295-331  gettop1000_nms
  * rerank_nms    eval_utils.py:230-260  gettopN_nms
Parity pinned: tests/golden/proposals.json holds the outputs of the reference's own functions on seeded
inputs (tests/golden/make_proposals_golden.py, run where /root/reference exists); tests/test_proposals.py
checks this file against them.  Equal scores: numpy's default argsort is unstable, so the order of exactly
equal scores is open in the reference; here (and on the GPU) the later proposal goes first, which is what a
stable sort gives.  The golden inputs have distinct scores.
"""
import numpy as np


def topn_select(scores, mask, topn=1000, val_score_thres=0):
    """eval_utils.py:263-279.  scores, mask [T,K].  Returns (ind, start, end, conf) arrays in the order of the
    reference's lists: cells (n,k), n >= k, with p >= max(topn-th largest p, val_score_thres), row-major."""
    p = np.asarray(scores) * np.asarray(mask)
    T, K = p.shape
    flat = np.sort(p.reshape(-1))
    thr = flat[-min(flat.size, topn)]
    good = p >= max(thr, val_score_thres)
    n_idx, k_idx = np.nonzero(good & (np.arange(T)[:, None] >= np.arange(K)[None, :]))  # row-major order
    return n_idx.astype(np.int64), (n_idx - k_idx).astype(np.int64), (n_idx + 1).astype(np.int64), p[n_idx, k_idx]


def enumerate_proposals(T, K):
    """eval_utils.py:305-310: (n,k) for n in range(T) for k in range(min(n,K)) -> arrays n, k."""
    n = np.repeat(np.arange(T), np.minimum(np.arange(T), K))
    first = np.cumsum(np.minimum(np.arange(T), K)) - np.minimum(np.arange(T), K)
    k = np.arange(n.size) - first[n]
    return n, k


def nms_select(scores, overlap=0.8, topn=1000):
    """eval_utils.py:295-325.  Returns (pick, props[P,2], nms_scores): pick = enumeration indices in pick order."""
    scores = np.asarray(scores)
    T, K = scores.shape
    n, k = enumerate_proposals(T, K)
    s = scores[n, k].astype(np.float64)
    t1 = (n - k).astype(np.int64)
    t2 = (n + 1).astype(np.int64)
    area = (t2 - t1 + 1).astype(float)
    order = np.argsort(s, kind="stable")[::-1]  # best first; equal scores: later proposal first
    alive = np.ones(s.size, dtype=bool)
    pick = []
    for i in order:
        if len(pick) >= topn:
            break
        if not alive[i]:
            continue
        pick.append(int(i))
        alive[i] = False
        wh = np.maximum(0., np.minimum(t2[i], t2) - np.maximum(t1[i], t1) + 1.0)
        o = wh / (area[i] + area - wh)
        alive &= (o <= overlap)
    pick = np.asarray(pick, dtype=np.int64)
    return pick, np.stack([t1[pick], t2[pick]], 1) if pick.size else np.zeros((0, 2), np.int64), s[pick]


def rerank_nms(props, prop_scores, sent_score, nms_overlap=0.999, topn=1000):
    """eval_utils.py:230-258.  Returns the pick list (indices into props)."""
    props = np.asarray(props, dtype=float)
    ps = np.asarray(prop_scores, dtype=float)
    ss = np.asarray(sent_score, dtype=float)
    t1, t2 = props[:, 0], props[:, 1]
    area = (t2 - t1 + 1e-3).astype(float)
    rank = np.empty(ps.size, dtype=np.int64)  # position in the ascending (stable) order of prop_scores
    rank[np.argsort(ps, kind="stable")] = np.arange(ps.size)
    alive = np.ones(ps.size, dtype=bool)
    pick = []
    while alive.any() and len(pick) < topn:
        cand = np.nonzero(alive)[0]
        i = cand[np.argmax(rank[cand])]  # ind[-1]
        wh = np.maximum(0., np.minimum(t2[i], t2) - np.maximum(t1[i], t1) + 1e-3)
        o = wh / (area[i] + area - wh)
        same = np.nonzero(alive & (o >= nms_overlap))[0]
        same = same[np.argsort(rank[same])]  # the reference scans `ind`, i.e. ascending prop_score order
        pick.append(int(same[np.argmax(ss[same])]))
        alive &= (o <= nms_overlap)
    return pick
