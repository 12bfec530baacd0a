"""SST proposal encoder.  CPU: the oracle restatement against the reference module's own
outputs (tests/golden/sst.npz).  GPU: csrc/sst.cu through echr_b200.sst_model.SST against both."""
import os

import numpy as np
import pytest
import torch

import sst_cases as C
from oracle import sst_oracle as SO

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "sst.npz"))
torch.set_grad_enabled(False)
TOL = {"fp32": 1e-4, "bf16w": 1e-4, "f16": 2e-3}


@pytest.mark.parametrize("c", range(len(C.CASE_LENGTHS)))
def test_oracle_matches_reference_module(c):
    tap, pred = SO.sst_forward(C.state_dict(), C.features(c))
    assert np.abs(tap.numpy() - G["tap%d" % c]).max() < 2e-6
    assert np.abs(pred.numpy() - G["pred%d" % c]).max() < 2e-6


def test_reference_checkpoint_keys_load_strictly():
    from echr_b200.sst_model import SST
    m = SST(C.Opt())
    m.load_state_dict(C.state_dict(), strict=True)
    assert sorted(m.state_dict().keys()) == sorted(C.state_dict().keys())


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["fp32", "bf16w", "f16"])
def test_gpu_forward_matches_reference(precision):
    from echr_b200.sst_model import SST
    sd = C.state_dict()
    if precision != "fp32":  # a bf16 run holds bf16-representable weights on both sides
        sd = {k: v.bfloat16().float() for k, v in sd.items()}
    m = SST(C.Opt(), precision=precision)
    m.load_state_dict(sd, strict=True)
    m = m.cuda()
    m.eval()
    feats = [C.features(c) for c in range(len(C.CASE_LENGTHS))]
    single = [m(f.cuda()) for f in feats]
    # the same shape three times: eager, captured into a graph, replayed — all three must agree bitwise
    runs = [[(a.clone(), b.clone()) for a, b in m.forward_batch([f.cuda() for f in feats])] for _ in range(3)]
    for r in runs[1:]:
        for (a0, b0), (a1, b1) in zip(runs[0], r):
            assert torch.equal(a0, a1) and torch.equal(b0, b1)
    batch = runs[2]
    worst = 0.0
    for c, f in enumerate(feats):
        if precision == "fp32":
            want_tap, want_pred = G["tap%d" % c], G["pred%d" % c]
        else:
            wt, wp = SO.sst_forward(sd, f)
            want_tap, want_pred = wt.numpy(), wp.numpy()
        for got_tap, got_pred in (single[c], batch[c]):
            assert tuple(got_tap.shape) == want_tap.shape and tuple(got_pred.shape) == want_pred.shape
            e1 = np.abs(got_tap.cpu().numpy() - want_tap).max()
            e2 = np.abs(got_pred.cpu().numpy() - want_pred).max()
            worst = max(worst, e1, e2)
            assert e1 < TOL[precision] and e2 < TOL[precision], (precision, c, e1, e2)
    print("sst %s max |d| = %.2e" % (precision, worst))


@pytest.mark.gpu
def test_gpu_ragged_batch_with_empty_video_and_many_videos():
    from echr_b200.sst_model import SST
    sd = C.state_dict(1, k=64)

    class Opt(C.Opt):
        K = 64
    m = SST(Opt(), precision="fp32")
    m.load_state_dict(sd, strict=True)
    m = m.cuda()
    m.eval()
    rng = np.random.RandomState(4)
    lens = [3, 0, 17, 1, 17, 9] + [int(x) for x in rng.randint(1, 30, size=150)]
    feats = [torch.from_numpy(rng.randn(n, 500).astype(np.float32)) for n in lens]
    got = m.forward_batch([f.cuda() for f in feats])
    for i in list(range(6)) + [20, 77, 155]:
        wt, wp = SO.sst_forward(sd, feats[i])
        assert tuple(got[i][0].shape) == (lens[i], 512) and tuple(got[i][1].shape) == (lens[i], 64)
        if lens[i]:
            assert np.abs(got[i][0].cpu().numpy() - wt.numpy()).max() < 1e-4
            assert np.abs(got[i][1].cpu().numpy() - wp.numpy()).max() < 1e-4


@pytest.mark.gpu
def test_gpu_sst_feeds_selection_and_caption_path():
    """The reference's eval flow on the device end to end (eval_utils.py:55,114-136): SST -> gettop1000 ->
    CaptionGenerator, no [T,K] matrix on the host."""
    import helpers
    import proposals_cases as PC
    from echr_b200 import proposals as P
    from echr_b200 import synth
    from echr_b200.caption_generator import CaptionGenerator
    from echr_b200.config import echr_opt
    from echr_b200.sst_model import SST
    from oracle import proposals_oracle as PO
    sst = SST(C.Opt(), precision="fp32")
    sst.load_state_dict(C.state_dict(), strict=True)
    sst = sst.cuda()
    sst.eval()
    v = synth.make_video(5, T=40, n_events=3)
    tap_feats, pred = sst(v.c3d_feats.cuda())
    mask = (np.arange(40)[:, None] >= np.arange(C.K)[None, :]).astype(np.float32)
    ind, feat, _, _, conf = P.gettop1000(pred, mask, [], 80.0, PC.featstamp_to_time, topN=10)
    wi, ws, we, _ = PO.topn_select(pred.cpu().numpy(), mask, topn=10)
    assert ind == wi.tolist() and feat == np.stack([ws, we], 1).tolist()
    cg = CaptionGenerator(echr_opt(), precision="fp32")
    cg.load_state_dict(helpers.state_dict("fp32"))
    cg = cg.cuda()
    seq, logp = cg(tap_feats, v.c3d_feats.cuda(), v.lda_feats.cuda(), [], ind, feat, mode="eval")
    assert seq.shape[0] == len(ind) and torch.isfinite(logp).all()
