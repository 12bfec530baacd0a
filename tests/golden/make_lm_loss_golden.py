"""Generates tests/golden/lm_loss.json: the validation LM loss exactly as eval_utils.py:139-152 computes it —
the UNMODIFIED reference model (through oracle/ref_shim.py) in mode='train' followed by the reference's own
misc.utils.LanguageModelCriterion — on the golden cases, with labels = BOS + the reference's greedy caption and
the loader's mask convention (1 up to and including the END token, dataloader.py mask_batch).

    python tests/golden/make_lm_loss_golden.py
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from echr_b200 import synth  # noqa: E402
from echr_b200.config import echr_opt  # noqa: E402
from oracle import ref_shim  # noqa: E402
from make_golden import CASES  # noqa: E402

sys.path.insert(0, "/root/reference")
import misc.utils as ref_utils  # noqa: E402  (the reference)


def labels_and_masks(seq, cols=32):
    """labels [N, cols] = [BOS, tokens..., 0...]; masks 1 for BOS, the tokens and the END that follows them."""
    n, s = seq.shape
    labels = torch.zeros(n, cols, dtype=torch.long)
    labels[:, 1:1 + s] = seq
    masks = torch.zeros(n, cols)
    for i in range(n):
        nz = int((seq[i] > 0).sum())
        masks[i, : nz + 2] = 1
    return labels, masks


def main():
    torch.set_grad_enabled(False)
    opt = echr_opt()
    out = {}
    crit = ref_utils.LanguageModelCriterion()
    for wname, rounder in (("fp32", None), ("bf16w", synth.round_state_dict_bf16)):
        sd = synth.init_state_dict(opt, seed=0)
        if rounder:
            sd = rounder(sd)
        model = ref_shim.build_reference(opt, sd)
        for cname, (seed, T, n, style) in CASES.items():
            v = synth.make_video(seed, T=T, n_events=n, style=style, bf16_features=(wname == "bf16w"))
            a = v.args()
            seq, _ = model(*a, mode="eval")
            labels, masks = labels_and_masks(seq)
            pred = model(a[0], a[1], a[2], labels, a[4], a[5], mode="train")
            loss = crit(pred, labels[:, 1:], masks[:, 1:])
            out["%s_%s" % (cname, wname)] = {"loss": float(loss), "steps": int(pred.shape[1])}
            print(cname, wname, float(loss), tuple(pred.shape))
    json.dump(out, open(os.path.join(ROOT, "tests", "golden", "lm_loss.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
