"""Output stage against golden vectors produced by the reference's own
misc/utils.decode_sequence and eval_utils.reranking (tests/golden/make_output_golden.py)."""
import json
import os

import numpy as np
import pytest
import torch

from echr_b200 import output_stage as OS

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "output_stage.json")))


@pytest.mark.parametrize("case", range(len(GOLD["cases"])))
def test_decode_sequence_and_entries_match_reference(case):
    c = GOLD["cases"][case]
    vocab = GOLD["vocab"]
    seq = torch.tensor(c["seq"], dtype=torch.long)
    assert OS.decode_sequence(vocab, seq) == c["sents"]
    assert OS.decode_sequence(OS.Vocab(vocab), np.asarray(c["seq"], dtype=np.int32)) == c["sents"]
    info = OS.video_entries(c["sents"], c["stamps"], c["score"], c["tap"])
    assert info == c["info"]
    assert OS.reranking(info) == c["reranked"]
    lp = torch.zeros(len(c["score"]), 30)
    lp[:, 0] = torch.tensor(c["score"])
    got = OS.caption_results(vocab, seq, lp, c["stamps"], c["tap"], is_reranking=True)
    assert [g["sentence"] for g in got] == [g["sentence"] for g in c["reranked"]]
    assert np.allclose([g["re_score"] for g in got], [g["re_score"] for g in c["reranked"]], atol=1e-4)


def test_edge_cases():
    vocab = {"1": "a", "2": "b"}
    assert OS.decode_sequence(vocab, []) == []                       # the decoder's empty result ([], [])
    assert OS.caption_results(vocab, [], [], [], []) == []
    assert OS.decode_sequence(vocab, torch.zeros(2, 5, dtype=torch.long)) == ["", ""]
    assert OS.decode_sequence(vocab, torch.tensor([[1, 2, 0, 2]])) == ["a b"]   # stops at the first END
    assert OS.decode_sequence(vocab, torch.tensor([[2, 2, 2]])) == ["b b b"]    # no END in the row
    with pytest.raises(KeyError):                                    # like ix_to_word[str(ix)] in the reference
        OS.decode_sequence(vocab, torch.tensor([[1, 7, 0]]))
    assert OS.decode_sequence(vocab, torch.tensor([[1, 0, 7]])) == ["a"]        # ids after END are never looked up
    assert OS.reranking([]) == []
    assert OS.sentence_scores([]).shape == (0,)
