"""The one model configuration this library implements: the ECHR recipe.

Reference: experiments/train_ECHR.sh:5 (flags) + opts.py:72-160 (defaults).  The reference threads one
mutable argparse namespace through every constructor (CaptionGenerator.py:56-84 and
models/MA_attention_8_NEW.py:13-22 write derived sizes back into it); `EchrOpt` is that namespace with
the derived fields already filled in, and `validate()` rejects every other recipe loudly — there is
no multi-variant dispatch here.
"""
from types import SimpleNamespace

ECHR_RECIPE = dict(
    fusion_model="TSRM8", event_context_type="ER3", video_context_type="VL", clip_context_type="CC",
    lda_dim=100, video_dim=500, hidden_dim=512, K=256,
    d_feats=512, d_o=512, n_head=16, use_posit=1, fST_type="fST0",
    caption_model="three_stream", CG_num_layers=3, CG_init_feats_type="", CG_input_feats_type="",
    CG_vocab_size=5747, CG_input_encoding_size=512, CG_rnn_type="lstm", CG_rnn_size=512,
    CG_drop_prob=0.5, CG_seq_length=30, CG_att_hid_size=512, CG_fc_feat_size=512,
    # upstream proposal encoder (models/sst_model.py), only used to size tap_feats
    tap_model="SST", tap_rnn_type="LSTM", rnn_num_layers=2, rnn_dropout=0.5,
)

# fields the kernels are specialised for (any other value is refused)
_FIXED = ("fusion_model", "event_context_type", "video_context_type", "clip_context_type", "video_dim",
          "hidden_dim", "lda_dim", "d_feats", "d_o", "n_head", "use_posit", "caption_model",
          "CG_num_layers", "CG_init_feats_type", "CG_input_feats_type", "CG_input_encoding_size",
          "CG_rnn_size", "CG_att_hid_size", "CG_rnn_type")


class EchrOpt(SimpleNamespace):
    def copy(self):
        return EchrOpt(**vars(self))


def echr_opt(**overrides):
    """Namespace with the ECHR recipe (overrides allowed for vocab size / seq length only)."""
    d = dict(ECHR_RECIPE)
    d.update(overrides)
    opt = EchrOpt(**d)
    derive(opt)
    return opt


def derive(opt):
    """Fill the fields the reference constructors write back (CaptionGenerator.py:56-84, MA:13-22)."""
    opt.video_context_dim = opt.lda_dim
    opt.event_context_dim = opt.d_o
    opt.clip_context_dim = opt.video_dim
    opt.TSRM_input_dim = opt.video_dim + opt.hidden_dim
    opt.d_pos_vec = opt.d_feats
    return opt


def validate(opt):
    for k in _FIXED:
        have = getattr(opt, k, ECHR_RECIPE[k] if k == "fST_type" else None)
        if have != ECHR_RECIPE[k]:
            raise ValueError("echr_b200 implements only the ECHR recipe: %s=%r is unsupported (need %r)"
                             % (k, have, ECHR_RECIPE[k]))
    if getattr(opt, "fST_type", "fST0") != "fST0":
        raise ValueError("echr_b200 implements only fST_type='fST0'")
    if not (1 <= opt.CG_seq_length <= 64):
        raise ValueError("CG_seq_length must be in 1..64")
    if not (1 <= opt.CG_vocab_size <= 65535):
        raise ValueError("CG_vocab_size must be in 1..65535")
    return derive(opt)
