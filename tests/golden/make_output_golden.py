"""Generates tests/golden/output_stage.json from the UNMODIFIED reference functions
(misc/utils.py:24-38 decode_sequence, eval_utils.py:334-345 reranking) in the authoring container.

    python tests/golden/make_output_golden.py
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, "/root/reference")
import eval_utils  # noqa: E402  (the reference)
import misc.utils as ref_utils  # noqa: E402


def main():
    rng = np.random.RandomState(7)
    vocab = {str(i): "w%d" % i for i in range(1, 60)}
    cases = []
    for n, d in ((1, 1), (3, 30), (7, 12), (12, 30), (5, 4)):
        seq = rng.randint(0, 60, size=(n, d))
        seq[rng.rand(n, d) < 0.15] = 0
        if n > 2:
            seq[1, 0] = 0          # empty sentence
            seq[2, :] = rng.randint(1, 60, size=d)  # no END inside the row
        sents = ref_utils.decode_sequence(vocab, torch.from_numpy(seq))
        tap = rng.rand(n).round(4).tolist()
        score = (-rng.rand(n) * 30).round(4).tolist()
        stamps = [[float(i), float(i + 1)] for i in range(n)]
        info = []
        for i, s in enumerate(sents):  # eval_utils.py:157-167
            info.append({"sentence": s, "timestamp": stamps[i], "sentence_confidence": score[i], "proposal_score": tap[i],
                         "re_score": 10 * tap[i] + score[i], "num": [i, len(sents)]})
        cases.append({"seq": seq.tolist(), "sents": sents, "tap": tap, "score": score, "stamps": stamps,
                      "info": info, "reranked": eval_utils.reranking(info)})
    json.dump({"vocab": vocab, "cases": cases}, open(os.path.join(ROOT, "tests", "golden", "output_stage.json"), "w"), indent=1)
    print("wrote", len(cases), "cases")


if __name__ == "__main__":
    main()
