"""Deterministic synthetic weights and ActivityNet-shaped inputs.

Nothing here is on the compute path: it only manufactures the checkpoint (`cg_model` state_dict in
the reference's layout, train.py:456-461) and the per-video inputs the hot path consumes
(eval_utils.py:52-66,107-115,129; dataloader.py:47-53,496-499,509-524).  All randomness comes from
torch's CPU generator, which is reproducible across machines for one torch version, so the GPU box
regenerates exactly what the golden fixtures were made from.
"""
import math

import numpy as np
import torch


def _uniform(gen, shape, bound):
    return (torch.rand(shape, generator=gen, dtype=torch.float32) * 2.0 - 1.0) * bound


def init_state_dict(opt, seed=0):
    """fp32 `cg_model` state_dict with the reference's key names/shapes and
    torch-default-like init ranges (embed/logit U(-0.1,0.1), logit bias 0: OldModel_NEW.py:66-70)."""
    g = torch.Generator().manual_seed(seed)
    H, E, V1 = opt.CG_rnn_size, opt.CG_input_encoding_size, opt.CG_vocab_size + 1
    D, NH = opt.d_feats, opt.n_head
    sd = {}

    def linear(name, out_f, in_f):
        b = 1.0 / math.sqrt(in_f)
        sd[name + ".weight"] = _uniform(g, (out_f, in_f), b)
        sd[name + ".bias"] = _uniform(g, (out_f,), b)

    linear("fusion_model.h2a_layer", 10, 10)  # dead parameter (MA:23), kept for strict loading
    linear("fusion_model.event_emb", D, opt.TSRM_input_dim)
    linear("fusion_model.enc_attn.pair_pos_fc1", D, D)
    linear("fusion_model.enc_attn.pair_pos_fc2", NH, D)
    linear("fusion_model.enc_attn.query_1", D, D)
    linear("fusion_model.enc_attn.key_1", D, D)
    b = 1.0 / math.sqrt(D)  # grouped 1x1 conv: fan_in = in_channels / groups
    sd["fusion_model.enc_attn.linear_out_1.weight"] = _uniform(g, (opt.d_o, D, 1, 1), b)
    sd["fusion_model.enc_attn.linear_out_1.bias"] = _uniform(g, (opt.d_o,), b)
    sd["lm_model.embed.weight"] = _uniform(g, (V1, E), 0.1)
    sd["lm_model.logit.weight"] = _uniform(g, (V1, 3 * H), 0.1)
    sd["lm_model.logit.bias"] = torch.zeros(V1)
    b = 1.0 / math.sqrt(H)
    for l, ctx in enumerate((opt.event_context_dim, opt.clip_context_dim, opt.video_context_dim)):
        p = "lm_model.core.layer%d." % l
        sd[p + "weight_ih"] = _uniform(g, (4 * H, E + ctx), b)
        sd[p + "weight_hh"] = _uniform(g, (4 * H, H), b)
        sd[p + "bias_ih"] = _uniform(g, (4 * H,), b)
        sd[p + "bias_hh"] = _uniform(g, (4 * H,), b)
    linear("lm_model.core.fusion_layer", H, 3 * H)  # dead parameter (OldModel_NEW.py:783)
    linear("lm_model.core.attention.ctx2att", opt.CG_att_hid_size, opt.clip_context_dim)
    linear("lm_model.core.attention.h2att", opt.CG_att_hid_size, H)
    linear("lm_model.core.attention.alpha_net", 1, opt.CG_att_hid_size)
    return sd


def round_bf16(t):
    """Round an fp32 tensor to bf16-representable fp32 values (used on BOTH sides of a bf16 run)."""
    return t.to(torch.bfloat16).to(torch.float32)


def round_state_dict_bf16(sd):
    return {k: round_bf16(v) for k, v in sd.items()}


class Video(object):
    """One video's inputs in the reference's calling convention (eval_utils.py:129)."""
    __slots__ = ("tap_feats", "c3d_feats", "lda_feats", "ind_select_list", "soi_select_list")

    def __init__(self, tap_feats, c3d_feats, lda_feats, ind_select_list, soi_select_list):
        self.tap_feats, self.c3d_feats, self.lda_feats = tap_feats, c3d_feats, lda_feats
        self.ind_select_list, self.soi_select_list = ind_select_list, soi_select_list

    @property
    def n_events(self):
        return len(self.soi_select_list)

    def args(self, device=None):
        mv = (lambda t: t.to(device)) if device is not None else (lambda t: t)
        return (mv(self.tap_feats), mv(self.c3d_feats), mv(self.lda_feats), [],
                list(self.ind_select_list), [list(s) for s in self.soi_select_list])


def video_events(seed, T=100, n_events=None, style="gt", K=256):
    """The proposals of synthetic video `seed` without its features: (ind_select_list, soi_select_list).  The
    event generator (numpy RandomState) is independent of the feature generator (torch), so a job can lay out
    its shards from the event counts of every video before any rank builds features."""
    rs = np.random.RandomState(1000 + seed)
    ind, soi = [], []
    if style == "gt":
        n = int(n_events) if n_events is not None else int(rs.randint(3, 6))
        for _ in range(n):
            s = int(rs.randint(0, T - 1))
            ln = int(rs.randint(1, T // 2 + 1))
            e = min(s + ln, T)
            soi.append([s, e])
            ind.append(e - 1)
    elif style == "sst":
        n = int(n_events)
        seen = set()
        while len(soi) < n:
            end = int(rs.randint(1, T))
            k = int(rs.randint(0, min(end + 1, K)))
            if (end, k) in seen:
                continue
            seen.add((end, k))
            soi.append([end - k, end + 1])
            ind.append(end)
    else:
        raise ValueError(style)
    return ind, soi


def make_video(seed, T=100, n_events=None, style="gt", K=256, bf16_features=False):
    """Synthetic video `seed`: c3d ~ N(0,1) (the loader standardises, dataloader.py:49-51), tap_feats
    ~ U(-1,1) (an LSTM output range, sst_model.py:36-38), lda = softmax(N(0,1)).
    style 'gt'  : n in {3,4,5} (or n_events), s~U{0..T-2}, len~U{1..T/2}, e=min(s+len,T), ind=e-1
                  (dataloader.py:496-499).
    style 'sst' : n_events distinct proposals ending at clip n with k clips before it:
                  soi=[n-k, n+1), ind=n, k < min(n+1, K) (eval_utils.py:278-284; K caps the span).
    """
    g = torch.Generator().manual_seed(1000 + seed)
    c3d = torch.randn((T, 500), generator=g, dtype=torch.float32)
    tap = torch.rand((T, 512), generator=g, dtype=torch.float32) * 2.0 - 1.0
    lda = torch.softmax(torch.randn((100,), generator=g, dtype=torch.float32), 0)
    if bf16_features:
        c3d, tap, lda = round_bf16(c3d), round_bf16(tap), round_bf16(lda)
    ind, soi = video_events(seed, T, n_events, style, K)
    return Video(tap, c3d, lda, ind, soi)


# BASELINE.json configs -> (n_videos, T, n_events, style, beam)
WORKLOADS = {
    "cfg1": dict(n_videos=1, T=100, n_events=3, style="gt", beam=1),
    "cfg2": dict(n_videos=64, T=100, n_events=None, style="gt", beam=1),
    "cfg3": dict(n_videos=32, T=100, n_events=100, style="sst", beam=3),
    "cfg4": dict(n_videos=16, T=400, n_events=200, style="sst", beam=1),
    "cfg5": dict(n_videos=4096, T=100, n_events=None, style="gt", beam=1),
}


def make_workload(name, n_videos=None, first_seed=0, bf16_features=False):
    w = dict(WORKLOADS[name])
    if n_videos is not None:
        w["n_videos"] = n_videos
    vids = [make_video(first_seed + v, T=w["T"], n_events=w["n_events"], style=w["style"],
                       bf16_features=bf16_features) for v in range(w["n_videos"])]
    return vids, w
