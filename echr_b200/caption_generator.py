"""Drop-in mirror of the reference's caption-generation call surface, running on the B200 engine.

Mirrors (same names, arguments, return contract, parameter names/shapes):
  CaptionGenerator            CaptionGenerator.py:7-53          forward(..., mode='eval'|'train')
  MA_Attention8 (fusion)      models/MA_attention_8_NEW.py:9-49
  ThreestreamModel (lm)       models/OldModel_NEW.py:18-187,1040-1043   forward / sample
  setup_lm / setup_fusion     models/__init__.py:6-29

The nn.Parameters here are only the checkpoint container (`load_state_dict(strict=True)` of a
reference `cg_model` works); all arithmetic runs in libechr_b200.so.  Additive API:
`caption_batch(videos, beam_size)` for many videos per call and `beam_size` in `forward(mode='eval')`.
"""
import weakref

import numpy as np
import torch
from torch import nn

from . import config as _config
from .engine import BatchLayout, Engine


class _EncAttn(nn.Module):  # attention_module_multi_head parameters (MA:82-100)
    def __init__(self, opt):
        super().__init__()
        d, g = opt.d_feats, opt.n_head
        self.pair_pos_fc1 = nn.Linear(opt.d_pos_vec, opt.d_pos_vec)
        self.pair_pos_fc2 = nn.Linear(opt.d_pos_vec, g)
        self.query_1 = nn.Linear(d, d)
        self.key_1 = nn.Linear(d, d)
        self.linear_out_1 = nn.Conv2d(g * d, opt.d_o, kernel_size=(1, 1), stride=1, groups=g)


class MA_Attention8(nn.Module):
    def __init__(self, opt):
        super().__init__()
        _config.validate(opt)
        self.h2a_layer = nn.Linear(10, 10)  # dead parameter of the reference (MA:23)
        self.event_emb = nn.Linear(opt.TSRM_input_dim, opt.d_feats)
        self.enc_attn = _EncAttn(opt)
        self.output_dim, self.use_posit, self.d_pos_vec, self.fST_type = opt.d_o, opt.use_posit, opt.d_pos_vec, "fST0"

    def forward(self, feats, soi_select_list):
        """feats [N,1012] (pooled clip feature | proposal state), soi_select_list [[s,e),...] -> [N,512]
        (models/MA_attention_8_NEW.py:35-49), computed by the engine's relation-encoder kernels."""
        return _owner_engine(self).relation_encode(feats, [list(s) for s in soi_select_list])

    # Host-side helpers of the reference kept for callers that use them directly (MA:51-79, numpy
    # float64 like the reference; the engine evaluates the same formulas on the device in fp64).
    @staticmethod
    def extract_position_matrix(bbox, nongt_dim=None):
        bbox = np.asarray(bbox)
        start, end = bbox[:, 0:1].astype(np.float64), bbox[:, 1:2].astype(np.float64)
        center = 0.5 * (start + end)
        length = (bbox[:, 1:2] - bbox[:, 0:1]).astype(np.float32)
        delta_center = np.maximum(np.abs((center - center.T) / length), 1e-3)
        delta_length = np.log(length.T / length)
        return np.stack([delta_center, delta_length], axis=2)

    @staticmethod
    def extract_position_embedding(position_mat, feat_dim, wave_length=10000):
        k = np.arange(feat_dim // 4, dtype=np.float64)
        div = np.power(float(wave_length), (4.0 / feat_dim) * k)
        z = (100.0 * position_mat)[..., None] / div
        emb = np.concatenate([np.sin(z), np.cos(z)], axis=3)
        return emb.reshape(position_mat.shape[0], position_mat.shape[1], feat_dim)


class _Attention(nn.Module):  # OldModel_NEW.py:366-374
    def __init__(self, opt):
        super().__init__()
        self.ctx2att = nn.Linear(opt.clip_context_dim, opt.CG_att_hid_size)
        self.h2att = nn.Linear(opt.CG_rnn_size, opt.CG_att_hid_size)
        self.alpha_net = nn.Linear(opt.CG_att_hid_size, 1)


class _ThreeStreamCore(nn.Module):  # OldModel_NEW.py:762-788
    def __init__(self, opt):
        super().__init__()
        e, h = opt.CG_input_encoding_size, opt.CG_rnn_size
        self.layer0 = nn.LSTMCell(opt.event_context_dim + e, h)
        self.layer1 = nn.LSTMCell(opt.clip_context_dim + e, h)
        self.layer2 = nn.LSTMCell(opt.video_context_dim + e, h)
        self.fusion_layer = nn.Linear(3 * h, h)  # dead parameter of the reference (OldModel_NEW.py:783)
        self.attention = _Attention(opt)


def _owner_engine(module):
    ref = getattr(module, "_owner", None)
    owner = ref() if ref is not None else None
    if owner is None:
        raise RuntimeError("this sub-module runs on the engine of the CaptionGenerator that owns it; "
                           "construct echr_b200.CaptionGenerator(opt) and use its .fusion_model / .lm_model")
    return owner.engine


def _clip_lengths(clip_mask):
    """The reference's clip_mask (CaptionGenerator.py:140-167) marks a prefix of each row."""
    m = clip_mask.detach().to("cpu") != 0
    lens = m.sum(1).to(torch.int64)
    pos = torch.arange(m.shape[1]).unsqueeze(0)
    if not bool(((pos < lens.unsqueeze(1)) == m).all()) or int(lens.min()) < 1:
        raise ValueError("clip_mask must mark a non-empty prefix of every row")
    return lens.tolist()


class ThreestreamModel(nn.Module):
    def __init__(self, opt):
        super().__init__()
        _config.validate(opt)
        self.opt = opt
        self.vocab_size, self.seq_length = opt.CG_vocab_size, opt.CG_seq_length
        self.num_layers, self.rnn_size = opt.CG_num_layers, opt.CG_rnn_size
        self.embed = nn.Embedding(self.vocab_size + 1, opt.CG_input_encoding_size)
        self.logit = nn.Linear(3 * opt.CG_rnn_size, self.vocab_size + 1)
        self.core = _ThreeStreamCore(opt)

    def init_hidden(self, video, event, clip):
        """CG_init_feats_type == '' -> zero state (OldModel_NEW.py:72-78)."""
        n = event.shape[0]
        z = torch.zeros(self.num_layers, n, self.rnn_size, dtype=torch.float32, device=event.device)
        return z, z.clone()

    def forward(self, video, event, clip, clip_mask, seq):
        """Teacher-forced log-probs [N, steps, V+1] (OldModel_NEW.py:98-130), forward only."""
        with torch.no_grad():
            eng = _owner_engine(self)
            eng.set_context(video, event, clip, _clip_lengths(clip_mask))
            return eng.teacher_forced(seq[:, :_teacher_forced_cols(seq)])

    def sample(self, video, event, clip, clip_mask, opt={}):
        """Greedy (default) or beam decode from ready-made contexts (OldModel_NEW.py:139-187).
        opt: {'sample_max': 1, 'beam_size': k}; multinomial sampling (sample_max=0) is only used by the
        reference's dead RL branch and is not implemented."""
        if not opt.get("sample_max", 1):
            raise ValueError("echr_b200 implements sample_max=1 (greedy / beam) only")
        beam = int(opt.get("beam_size", 1))
        with torch.no_grad():
            eng = _owner_engine(self)
            eng.set_context(video, event, clip, _clip_lengths(clip_mask), beam=max(1, beam))
            if beam > 1:
                tok, lp, score = eng.decode_beam(beam)
                return tok.long(), lp
            tok, lp = eng.decode_greedy()
            return trim_greedy(tok, lp)

    def get_logprobs_state(self, it, video, event, clip, clip_mask, state):
        """One step with explicit state (OldModel_NEW.py:133-137): returns (logprobs, (h, c))."""
        with torch.no_grad():
            eng = _owner_engine(self)
            eng.set_context(video, event, clip, _clip_lengths(clip_mask))
            lp, h, c = eng.decode_step(it, state[0], state[1])
            return lp, (h, c)


def setup_lm(opt):
    if opt.caption_model != "three_stream":
        raise ValueError("echr_b200 implements caption_model='three_stream' only")
    return ThreestreamModel(opt)


def setup_fusion(opt):
    if opt.fusion_model != "TSRM8":
        raise ValueError("fusion model not supported: {}".format(opt.fusion_model))
    return MA_Attention8(opt)


def trim_greedy(tokens, logp):
    """The reference stops (and so trims the columns) at the first step where every row of the call
    has already produced token 0 (OldModel_NEW.py:175-183); an empty result is ([], [])."""
    alive = (tokens.cpu() != 0).any(dim=0).numpy()  # one small D2H (the call's only sync), the rest on the host
    dead = np.flatnonzero(~alive)
    ncols = int(dead[0]) if dead.size else tokens.shape[1]
    if ncols == 0:
        return [], []
    return tokens[:, :ncols].long(), logp[:, :ncols]


def _teacher_forced_cols(seq):
    """OldModel.forward (OldModel_NEW.py:98-130) stops at the first all-zero label column i >= 1: the columns
    of `seq` it consumes (it then returns that many minus one steps).  One reduction and one device-to-host
    copy instead of the reference's `.sum()` test per step."""
    cols = seq.shape[1]
    if cols > 2:
        sums = seq[:, 1:cols - 1].sum(0).tolist()
        for i, v in enumerate(sums, start=1):
            if v == 0:
                return i + 1
    return cols


class CaptionGenerator(nn.Module):
    def __init__(self, opt, precision="fp32"):
        super().__init__()
        self.opt = _config.validate(opt)
        self.precision = precision
        self.fusion_model = setup_fusion(opt)
        self.lm_model = setup_lm(opt)
        for sub in (self.fusion_model, self.lm_model):  # the sub-modules run on this module's engine
            object.__setattr__(sub, "_owner", weakref.ref(self))
        self._engine = None
        self._engine_versions = None
        self.eval()

    # -- engine lifetime: rebuilt whenever the parameters may have changed
    def _invalidate(self):
        if self._engine is not None:
            self._engine.close()
        self._engine = None

    def load_state_dict(self, state_dict, strict=True, **kw):
        out = super().load_state_dict(state_dict, strict=strict, **kw)
        self._invalidate()
        return out

    def _apply(self, fn, *a, **kw):
        out = super()._apply(fn, *a, **kw)
        self._invalidate()
        return out

    def refresh(self):
        """Re-pack the weights on the next call.  Needed only after edits torch cannot see: writes through
        `param.data` (they bypass the tensor version counters the engine watches)."""
        self._invalidate()

    def set_precision(self, precision):
        self.precision = precision
        self._invalidate()

    def _param_versions(self):
        # torch bumps a tensor's _version on every in-place write (optimizer steps, p.data.copy_, sub-module
        # load_state_dict): the packed engine is stale as soon as any of them moves
        return tuple(p._version for p in self.parameters())

    @property
    def engine(self):
        if self._engine is not None and self._engine_versions != self._param_versions():
            self._invalidate()
        if self._engine is None:
            dev = next(self.parameters()).device
            if dev.type != "cuda":
                raise RuntimeError("echr_b200.CaptionGenerator has no CPU path: call .cuda() first")
            self._engine = Engine(self.state_dict(), self.opt.CG_vocab_size, self.opt.CG_seq_length,
                                  precision=self.precision, device=dev)
            self._engine_versions = self._param_versions()
        return self._engine

    # -- batched API (additive)
    def _gather(self, videos):
        """Batch layout + the three feature tensors of a list of videos, on the engine's device."""
        eng = self.engine
        layout = BatchLayout([v[1].shape[0] for v in videos], [v[3] for v in videos], [v[4] for v in videos])
        f32 = lambda t: t.to(device=eng.device, dtype=torch.float32)
        if len(videos) == 1:  # the reference's call shape: no concatenation copies
            v = videos[0]
            c3d, tap, lda = f32(v[1]).contiguous(), f32(v[0]).contiguous(), f32(v[2]).reshape(1, -1).contiguous()
        else:
            c3d = torch.cat([f32(v[1]) for v in videos], 0).contiguous()
            tap = torch.cat([f32(v[0]) for v in videos], 0).contiguous()
            lda = torch.stack([f32(v[2]).reshape(-1) for v in videos], 0).contiguous()
        return eng, layout, c3d, tap, lda

    def _stage(self, videos, beam):
        eng, layout, c3d, tap, lda = self._gather(videos)
        eng.set_inputs(layout, c3d, tap, lda, beam=beam)
        eng.encode()
        return eng, layout

    def caption_batch(self, videos, beam_size=1):
        """videos: list of (tap_feats[T,512], c3d_feats[T,500], lda_feats[100], ind_select_list,
        soi_select_list).  Returns one (seq, logp) per video with the reference's per-call trimming
        (greedy) or (seq[N,S], logp[N,S], score[N]) per video (beam_size > 1)."""
        with torch.no_grad():
            eng, layout, c3d, tap, lda = self._gather(videos)
            if layout.n_events == 0:
                return [([], []) if beam_size <= 1 else (torch.zeros(0, eng.seq_length, dtype=torch.long, device=eng.device),
                                                        torch.zeros(0, eng.seq_length, device=eng.device),
                                                        torch.zeros(0, device=eng.device)) for _ in videos]
            tokens, logp, score = eng.caption_device(layout, c3d, tap, lda, beam=max(1, beam_size))  # one library call
            out, r0 = [], 0
            for n in layout.events_per_video:
                if beam_size > 1:
                    out.append((tokens[r0:r0 + n].long(), logp[r0:r0 + n], score[r0:r0 + n]))
                elif n == 0:
                    out.append(([], []))
                else:
                    out.append(trim_greedy(tokens[r0:r0 + n], logp[r0:r0 + n]))
                r0 += n
            return out

    def teacher_forced(self, videos, seqs):
        """Per-video teacher-forced log-probs (the `mode='train'` forward), no trimming."""
        with torch.no_grad():
            eng, layout = self._stage(videos, 1)
            seq = torch.cat([s.to(eng.device) for s in seqs], 0)
            lp = eng.teacher_forced(seq)
            out, r0 = [], 0
            for n in layout.events_per_video:
                out.append(lp[r0:r0 + n])
                r0 += n
            return out

    def lm_loss(self, tap_feats, c3d_feats, lda_feats, lm_labels, lm_masks, ind_select_list, soi_select_list):
        """The validation LM loss of eval_utils.py:139-152 in one pass:
        `cg_crit(cg_model(..., lm_labels, ..., mode='train'), lm_labels[:, 1:], lm_masks[:, 1:])` with
        cg_crit = LanguageModelCriterion (misc/utils.py:62-75), without the [N, S-1, V+1] log-prob tensor."""
        from .criterion import LanguageModelCriterion
        video = (tap_feats, c3d_feats, lda_feats, list(ind_select_list), [list(s) for s in soi_select_list])
        with torch.no_grad():
            eng, _ = self._stage([video], 1)
            cols = _teacher_forced_cols(lm_labels)
            tlp = eng.teacher_forced_nll(lm_labels[:, :cols])
            return LanguageModelCriterion.from_target_logp(tlp, lm_masks[:, 1:].to(tlp.device))

    # -- the reference's call (eval_utils.py:129, train.py:298)
    def forward(self, tap_feats, c3d_feats, lda_feats, lm_labels, ind_select_list, soi_select_list, mode="train",
                beam_size=1):
        video = (tap_feats, c3d_feats, lda_feats, list(ind_select_list), [list(s) for s in soi_select_list])
        if mode == "eval":
            res = self.caption_batch([video], beam_size)[0]
            return res if beam_size > 1 else (res[0], res[1])
        if mode == "train":
            # OldModel.forward (OldModel_NEW.py:98-130): stop at the first all-zero label column i >= 1
            seq = lm_labels
            return self.teacher_forced([video], [seq[:, :_teacher_forced_cols(seq)]])[0]
        raise ValueError("echr_b200.CaptionGenerator supports mode='eval' and mode='train' (forward only); got %r" % mode)
