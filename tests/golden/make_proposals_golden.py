"""Generates tests/golden/proposals.npz: the outputs of the UNMODIFIED reference functions gettop1000,
gettop1000_nms and gettopN_nms (eval_utils.py:230-331) on the seeded inputs of tests/proposals_cases.py.
Run in the authoring container (needs /root/reference):

    python tests/golden/make_proposals_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, "/root/reference")
import eval_utils  # noqa: E402  (the reference)
import proposals_cases as C  # noqa: E402


def main():
    out = {}
    for c in range(len(C.TOPN_CASES)):
        s, m, gts, topn, thres = C.topn_inputs(c)
        ind, feat, cg, ts, conf = eval_utils.gettop1000(s, m, gts, 120.0, C.featstamp_to_time, val_score_thres=thres, topN=topn)
        pre = "topn%d_" % c
        out[pre + "ind"] = np.asarray(ind, dtype=np.int64)
        out[pre + "feat"] = np.asarray(feat, dtype=np.int64).reshape(-1, 2)
        out[pre + "cg"] = np.asarray(cg, dtype=np.int64)
        out[pre + "ts"] = np.asarray(ts, dtype=np.float64).reshape(-1, 2)
        out[pre + "conf"] = np.asarray(conf, dtype=np.float64)
    for c in range(len(C.NMS_CASES)):
        s, gts, ov, topn = C.nms_inputs(c)
        ind, props, pg, ts, sc = eval_utils.gettop1000_nms(s, None, gts, 90.0, C.featstamp_to_time, overlap=ov, topN=topn)
        pre = "nms%d_" % c
        out[pre + "ind"] = np.asarray(ind, dtype=np.int64)
        out[pre + "props"] = np.asarray(props, dtype=np.int64).reshape(-1, 2)
        out[pre + "pg"] = np.asarray(pg, dtype=np.int64)
        out[pre + "ts"] = np.asarray(ts, dtype=np.float64).reshape(-1, 2)
        out[pre + "sc"] = np.asarray(sc, dtype=np.float64)
    for c in range(len(C.RERANK_CASES)):
        props, ps, ss, ov, topn = C.rerank_inputs(c)
        _, _, pick = eval_utils.gettopN_nms(props.tolist(), ps.tolist(), ss.tolist(), nms_overlap=ov, topN=topn)
        out["rr%d_pick" % c] = np.asarray(pick, dtype=np.int64)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "proposals.npz"), **out)
    print("wrote", {k: v.shape for k, v in out.items() if k.endswith(("ind", "pick"))})


if __name__ == "__main__":
    main()
