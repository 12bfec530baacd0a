"""Validation LM loss (eval_utils.py:139-152): the teacher-forced pass + LanguageModelCriterion
(misc/utils.py:62-75).  Golden = the reference model + the reference criterion (tests/golden/lm_loss.json).
CPU: oracle + criterion mirror against it.  GPU: CaptionGenerator.lm_loss (echr_teacher_forced_nll, no
[N,S,V+1] tensor) and the mirror criterion on the full mode='train' output."""
import json
import os

import numpy as np
import pytest
import torch

import helpers
from echr_b200.criterion import LanguageModelCriterion
from oracle import echr_oracle as O

G = json.load(open(os.path.join(helpers.GOLDEN_DIR, "lm_loss.json")))
torch.set_grad_enabled(False)


def labels_and_masks(seq, cols=32):
    seq = torch.as_tensor(np.asarray(seq)).long()
    n, s = seq.shape
    labels = torch.zeros(n, cols, dtype=torch.long)
    labels[:, 1:1 + s] = seq
    masks = torch.zeros(n, cols)
    for i in range(n):
        masks[i, : int((seq[i] > 0).sum()) + 2] = 1
    return labels, masks


def test_criterion_mirror_matches_the_formula_and_truncates():
    g = torch.Generator().manual_seed(0)
    lp = torch.log_softmax(torch.randn(5, 7, 11, generator=g), 2)
    target = torch.randint(0, 11, (5, 9), generator=g)
    mask = (torch.rand(5, 9, generator=g) > 0.3).float()
    want = -(lp.gather(2, target[:, :7].unsqueeze(2)).squeeze(2) * mask[:, :7]).sum() / (mask[:, :7].sum() + 1e-6)
    got = LanguageModelCriterion()(lp, target, mask)
    assert abs(float(got) - float(want)) < 1e-7
    assert float(LanguageModelCriterion()(lp, target, torch.zeros(5, 9))) == 0.0


@pytest.mark.parametrize("kind", ["fp32", "bf16w"])
@pytest.mark.parametrize("name", ["cfg1", "gt5"])
def test_oracle_loss_matches_reference(name, kind):
    g, v = helpers.golden(name, kind), helpers.case_video(name, kind)
    sd = helpers.state_dict(kind)
    labels, masks = labels_and_masks(g["seq"])
    video, event, clip, mask = O.build_context(sd, v.tap_feats, v.c3d_feats, v.lda_feats, v.ind_select_list, v.soi_select_list)
    lp = O.teacher_forced_logprobs(sd, video, event, clip, mask, labels)
    loss = LanguageModelCriterion()(lp, labels[:, 1:], masks[:, 1:])
    assert lp.shape[1] == G["%s_%s" % (name, kind)]["steps"]
    assert abs(float(loss) - G["%s_%s" % (name, kind)]["loss"]) < 2e-5


@pytest.mark.gpu
@pytest.mark.parametrize("precision,kind,tol", [("fp32", "fp32", 2e-5), ("bf16w", "bf16w", 2e-5), ("f16", "bf16w", 5e-4)])
@pytest.mark.parametrize("name", ["cfg1", "gt5", "sst12", "long30"])
def test_gpu_lm_loss_matches_reference(name, precision, kind, tol):
    from echr_b200.caption_generator import CaptionGenerator
    from echr_b200.config import echr_opt
    g, v = helpers.golden(name, kind), helpers.case_video(name, kind)
    m = CaptionGenerator(echr_opt(), precision=precision)
    m.load_state_dict(helpers.state_dict(kind), strict=True)
    m = m.cuda()
    labels, masks = labels_and_masks(g["seq"])
    vt = helpers.video_tuple(v, "cuda:0")
    want = G["%s_%s" % (name, kind)]["loss"]
    fused = m.lm_loss(vt[0], vt[1], vt[2], labels.cuda(), masks.cuda(), vt[3], vt[4])
    full = m(vt[0], vt[1], vt[2], labels.cuda(), vt[3], vt[4], mode="train")
    assert full.shape[1] == G["%s_%s" % (name, kind)]["steps"]
    unfused = LanguageModelCriterion()(full, labels[:, 1:].cuda(), masks[:, 1:].cuda())
    assert abs(float(fused) - want) < tol, (float(fused), want)
    assert abs(float(unfused) - want) < tol
    # the fused pass is the same arithmetic: per-step entries equal the gathered full tensor
    tlp = m.engine.teacher_forced_nll(labels.cuda())
    picked = full.gather(2, labels[:, 1:].cuda().unsqueeze(2)).squeeze(2)
    assert torch.equal(tlp, picked)
    # ragged labels: a caption cut to 5 tokens stops the pass early (OldModel_NEW.py:122) on both paths
    short = labels.clone()
    short[:, 6:] = 0
    smask = masks.clone()
    smask[:, 7:] = 0
    f2 = m.lm_loss(vt[0], vt[1], vt[2], short.cuda(), smask.cuda(), vt[3], vt[4])
    full2 = m(vt[0], vt[1], vt[2], short.cuda(), vt[3], vt[4], mode="train")
    assert full2.shape[1] == 6
    u2 = LanguageModelCriterion()(full2, short[:, 1:].cuda(), smask[:, 1:].cuda())
    assert abs(float(f2) - float(u2)) < 1e-6


def test_teacher_forced_column_rule_matches_the_reference_loop():
    """OldModel_NEW.py:104-123: the step loop breaks at the first i >= 1 whose label column sums to zero."""
    from echr_b200.caption_generator import _teacher_forced_cols

    def reference_steps(seq):
        steps = 0
        for i in range(seq.size(1) - 1):
            if i >= 1 and seq[:, i].sum() == 0:
                break
            steps += 1
        return steps
    g = torch.Generator().manual_seed(3)
    for cols in (2, 3, 5, 17, 32):
        for cut in (None, 1, 2, cols - 2, cols - 1):
            seq = torch.randint(1, 50, (4, cols), generator=g)
            seq[:, 0] = 0
            if cut is not None and 0 < cut < cols:
                seq[:, cut:] = 0
            assert _teacher_forced_cols(seq) - 1 == reference_steps(seq), (cols, cut)
    ragged = torch.tensor([[0, 5, 0, 0, 0], [0, 7, 8, 0, 0]])
    assert _teacher_forced_cols(ragged) - 1 == reference_steps(ragged) == 3
