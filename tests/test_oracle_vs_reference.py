"""Authoring-container test: the oracle against the live reference modules (skipped where
/root/reference does not exist — there the golden vectors in tests/golden/ carry the pin)."""
import pytest
import torch

import helpers
from echr_b200 import synth
from echr_b200.config import echr_opt
from oracle import echr_oracle as O
from oracle import ref_shim

pytestmark = pytest.mark.skipif(not ref_shim.available(), reason="reference checkout not present")
torch.set_grad_enabled(False)


@pytest.fixture(scope="module")
def ref_model():
    return ref_shim.build_reference(echr_opt(), helpers.state_dict("fp32"))


@pytest.mark.parametrize("seed,n,style", [(11, 4, "gt"), (12, 9, "sst")])
def test_live_differential(ref_model, seed, n, style):
    sd = helpers.state_dict("fp32")
    v = synth.make_video(seed, n_events=n, style=style)
    seq, lp = ref_model(*v.args(), mode="eval")
    s2, l2 = O.caption_video(sd, v.tap_feats, v.c3d_feats, v.lda_feats, v.ind_select_list, v.soi_select_list)
    assert torch.equal(seq, s2)
    assert float((lp - l2).abs().max()) < 2e-5
    a = v.args()
    ev_ref = ref_model.get_event_context(a[0], a[1], a[2], a[4], a[5])
    _, ev, _, _ = O.build_context(sd, v.tap_feats, v.c3d_feats, v.lda_feats, v.ind_select_list, v.soi_select_list)
    assert float((ev - ev_ref).abs().max()) < 2e-6


def test_reference_state_dict_layout(ref_model):
    """The synthetic checkpoint has exactly the reference's keys and shapes (strict load both ways)."""
    sd = helpers.state_dict("fp32")
    ref_sd = ref_model.state_dict()
    assert set(sd) == set(ref_sd)
    for k in sd:
        assert tuple(sd[k].shape) == tuple(ref_sd[k].shape), k
    from echr_b200.caption_generator import CaptionGenerator
    mine = CaptionGenerator(echr_opt())
    mine.load_state_dict(ref_sd, strict=True)
    assert set(mine.state_dict()) == set(ref_sd)
