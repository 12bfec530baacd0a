"""Seeded inputs of the proposal-selection parity cases, shared by the golden generator
(tests/golden/make_proposals_golden.py, which runs the reference on them) and tests/test_proposals*.py.
numpy's legacy RandomState stream is stable across versions, so the fixture only stores the outputs."""
import numpy as np

TOPN_CASES = [  # (T, K, topN, val_score_thres, mask density)
    (100, 256, 1000, 0, 1.0), (100, 256, 50, 0, 1.0), (30, 16, 1000, 0.5, 0.8), (1, 8, 1000, 0, 1.0),
    (64, 64, 1000, 0, 0.05), (400, 256, 1000, 0.3, 1.0), (17, 5, 7, 0, 1.0),
]
NMS_CASES = [  # (T, K, overlap, topN)
    (100, 256, 0.8, 1000), (100, 256, 0.5, 20), (30, 16, 0.8, 1000), (2, 8, 0.8, 1000), (300, 64, 0.9, 1000),
    (40, 256, 0.0, 1000), (400, 256, 0.7, 100),
]
RERANK_CASES = [(1, 0.999, 1000), (12, 0.999, 1000), (200, 0.7, 1000), (200, 0.3, 5), (1000, 0.5, 1000), (60, 0.999, 1000)]


def featstamp_to_time(start_f, end_f, nfeats, duration):
    """dataloader.py:298-302 without the loader object."""
    time_per_feat = duration / nfeats
    start = min(max(0, start_f * time_per_feat), duration - time_per_feat)
    end = max(end_f * time_per_feat, start + time_per_feat)
    return start, end


def distinct_scores(rng, T, K):
    """float32 scores in (0,1), all different (the reference's argsort leaves equal scores unordered)."""
    v = (np.arange(T * K, dtype=np.float64) + 0.5) / (T * K)
    v = v ** rng.uniform(0.5, 2.0)
    s = rng.permutation(v).astype(np.float32).reshape(T, K)
    assert np.unique(s).size == s.size
    return s


def topn_inputs(c):
    T, K, topn, thres, dens = TOPN_CASES[c]
    rng = np.random.RandomState(1000 + c)
    s = distinct_scores(rng, T, K)
    m = (np.arange(T)[:, None] >= np.arange(K)[None, :]).astype(np.float32)
    if dens < 1.0:
        m *= (rng.rand(T, K) < dens).astype(np.float32)
    gts = rng.randint(0, 5, size=(T, K))
    return s, m, gts, topn, thres


def nms_inputs(c):
    T, K, ov, topn = NMS_CASES[c]
    rng = np.random.RandomState(2000 + c)
    return distinct_scores(rng, T, K), rng.randint(0, 5, size=(T, K)), ov, topn


def rerank_inputs(c):
    n, ov, topn = RERANK_CASES[c]
    rng = np.random.RandomState(3000 + c)
    start = rng.uniform(0, 100, size=n)
    length = rng.uniform(0.5, 40, size=n)
    props = np.stack([start, start + length], 1)
    if n >= 12:  # exact duplicates and near-duplicates: clusters with more than one member
        props[n // 2:n // 2 + 4] = props[0]
        props[n // 3] = props[1] + 1e-4
    ps = rng.permutation((np.arange(n) + 0.5) / n)
    ss = -rng.permutation(np.arange(n) + 1.0)
    return props, ps, ss, ov, topn
