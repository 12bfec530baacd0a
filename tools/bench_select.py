"""Measurement of the proposal-selection row: 64 videos x T=100 x K=256 score matrices resident
in HBM -> selected records, CUDA events around 20 launches; the oracle port and the reference-shaped per-video
Python loop are timed on the host beside it.  Prints one JSON line.
usage: python tools/bench_select.py [n_videos] [T] [K]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch
from echr_b200 import proposals as P
from oracle import proposals_oracle as PO

V = int(sys.argv[1]) if len(sys.argv) > 1 else 64
T = int(sys.argv[2]) if len(sys.argv) > 2 else 100
K = int(sys.argv[3]) if len(sys.argv) > 3 else 256
rng = np.random.RandomState(0)
s = rng.rand(V * T, K).astype(np.float32)
m = np.tile((np.arange(T)[:, None] >= np.arange(K)[None, :]).astype(np.float32), (V, 1))
off = np.arange(V + 1) * T
d_s, d_m = torch.from_numpy(s).cuda(), torch.from_numpy(m).cuda()


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


ms_topn = timed(lambda: P.select_topn(d_s, d_m, off, topN=1000))
ms_nms = timed(lambda: P.select_nms(d_s, off, overlap=0.8, topN=1000))
t0 = time.perf_counter()
for v in range(min(V, 8)):
    PO.topn_select(s[off[v]:off[v + 1]], m[off[v]:off[v + 1]], topn=1000)
cpu_topn = (time.perf_counter() - t0) / min(V, 8) * V * 1e3
t0 = time.perf_counter()
for v in range(min(V, 4)):
    PO.nms_select(s[off[v]:off[v + 1]], overlap=0.8, topn=1000)
cpu_nms = (time.perf_counter() - t0) / min(V, 4) * V * 1e3
bytes_topn = s.nbytes + m.nbytes  # algorithmic: every score and mask cell read once
print(json.dumps({"row": "N2 proposal selection", "videos": V, "T": T, "K": K,
                  "gpu_ms": {"select_topn": ms_topn, "select_nms": ms_nms},
                  "cpu_oracle_ms": {"topn_select": cpu_topn, "nms_select": cpu_nms, "cores": 1,
                                    "sample": "first 8 / 4 videos, scaled to %d" % V},
                  "topn_algorithmic_GBs": bytes_topn / (ms_topn * 1e-3) / 1e9,
                  "note": "select_topn re-reads the matrix 5x (4 radix passes + compaction), L2-resident; one CTA per video"}))
