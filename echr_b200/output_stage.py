"""Output stage after the decoder: token ids -> sentences -> per-video result entries.

Mirrors, on the host side of the boundary, what the reference does after `cg_model(...)` returns:
  * `decode_sequence`   misc/utils.py:24-38   (one `.item()` per token there; one D2H + a table lookup here)
  * `sentence_scores`   eval_utils.py:134     (`cg_prob.sum(1).cpu().numpy().astype('float')`)
  * `video_entries`     eval_utils.py:157-167 (the json schema of `tap_cg_pred[vid]`)
  * `reranking`         eval_utils.py:334-345 (keep the entries whose re_score reaches the 10th best)
Nothing here runs on the GPU: the decoder already returns at most [events, 30] ids per call.
"""
import numpy as np
import torch


class Vocab(object):
    """ix_to_word of the loader (`loader.get_vocab()`: {"1": "a", ...}) as an array for vectorised lookup."""

    def __init__(self, ix_to_word):
        top = max(int(k) for k in ix_to_word) if len(ix_to_word) else 0
        self.words = np.empty(top + 1, dtype=object)
        self.words[:] = ""
        self.known = np.zeros(top + 1, dtype=bool)
        for k, w in ix_to_word.items():
            self.words[int(k)] = w
            self.known[int(k)] = True
        self.ix_to_word = ix_to_word


def _as_numpy_ids(seq):
    if isinstance(seq, torch.Tensor):
        return seq.detach().to("cpu").numpy()  # the only device -> host copy of the stage
    return np.asarray(seq)


def decode_sequence(ix_to_word, seq):
    """misc/utils.py:24-38.  seq: [N, D] ids (tensor on any device, or array), 0 = END.  Returns N strings:
    the words of each row up to (not including) its first id <= 0, joined by single blanks."""
    if isinstance(seq, (list, tuple)) and len(seq) == 0:
        return []
    vocab = ix_to_word if isinstance(ix_to_word, Vocab) else Vocab(ix_to_word)
    ids = _as_numpy_ids(seq)
    if ids.ndim != 2:
        raise ValueError("decode_sequence expects a [N, D] id matrix, got shape %s" % (ids.shape,))
    n, d = ids.shape
    stop = ids <= 0
    length = np.where(stop.any(axis=1), stop.argmax(axis=1), d)  # first END per row
    # the reference raises KeyError on a (used) id outside the vocabulary
    used = np.arange(d)[None, :] < length[:, None]
    if used.any():
        uid = ids[used]
        if int(uid.max()) >= len(vocab.words) or not vocab.known[uid].all():
            bad = [int(x) for x in uid if int(x) >= len(vocab.words) or not vocab.known[int(x)]]
            raise KeyError(str(bad[0]))
    words = vocab.words[np.clip(ids, 0, len(vocab.words) - 1)]
    return [" ".join(words[i, : length[i]]) for i in range(n)]


def sentence_scores(cg_prob):
    """eval_utils.py:134: per-event sum of the step log-probs as float64 numpy."""
    if isinstance(cg_prob, (list, tuple)) and len(cg_prob) == 0:
        return np.zeros(0, dtype="float")
    if isinstance(cg_prob, torch.Tensor):
        return cg_prob.sum(1).cpu().numpy().astype("float")
    return np.asarray(cg_prob).sum(1).astype("float")


def video_entries(sents, timestamps, cg_score, tap_prob):
    """eval_utils.py:157-167: one dict per proposal, in proposal order."""
    out = []
    for i, sent in enumerate(sents):
        out.append({
            "sentence": sent,
            "timestamp": timestamps[i],
            "sentence_confidence": cg_score[i],
            "proposal_score": tap_prob[i],
            "re_score": 10 * tap_prob[i] + cg_score[i],
            "num": [i, len(sents)],
        })
    return out


def reranking(vid_info):
    """eval_utils.py:334-345: keep the proposals whose re_score is at least the 10th largest one."""
    if len(vid_info) == 0:
        return []
    scores = np.sort(np.array([v["re_score"] for v in vid_info]))
    threshold = scores[-min(len(scores), 10)]
    return [v for v in vid_info if v["re_score"] >= threshold]


def caption_results(ix_to_word, seq, cg_prob, timestamps, tap_prob, is_reranking=False):
    """The whole stage for one video: what eval_utils.py:131-172 stores under tap_cg_pred[vid]
    (an empty list when the decoder produced nothing, eval_utils.py:131-132)."""
    if isinstance(seq, (list, tuple)) and len(seq) == 0:
        return []
    sents = decode_sequence(ix_to_word, seq)
    info = video_entries(sents, timestamps, sentence_scores(cg_prob), tap_prob)
    return reranking(info) if is_reranking else info
