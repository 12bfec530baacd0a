"""Seeded weights and inputs of the SST parity cases, shared by tests/golden/make_sst_golden.py (which runs the
reference's SST module on them) and tests/test_sst*.py."""
import numpy as np
import torch

K = 256
CASE_LENGTHS = [7, 40, 1, 100]  # clips per video


def state_dict(seed=0, k=K):
    """tap_model state_dict with nn.LSTM-scale values (uniform +-1/sqrt(512)), from numpy's stable stream."""
    rng = np.random.RandomState(100 + seed)
    u = lambda *shape: torch.from_numpy(((rng.rand(*shape) * 2 - 1) / np.sqrt(512.0)).astype(np.float32))
    sd = {}
    for l, in_f in enumerate((500, 512)):
        sd["rnn.weight_ih_l%d" % l] = u(2048, in_f)
        sd["rnn.weight_hh_l%d" % l] = u(2048, 512)
        sd["rnn.bias_ih_l%d" % l] = u(2048)
        sd["rnn.bias_hh_l%d" % l] = u(2048)
    sd["scores.weight"] = u(k, 512) * 4
    sd["scores.bias"] = u(k)
    return sd


def features(case):
    rng = np.random.RandomState(200 + case)
    return torch.from_numpy(rng.randn(CASE_LENGTHS[case], 500).astype(np.float32))


class Opt(object):
    hidden_dim, K, video_dim, tap_rnn_type, rnn_num_layers, rnn_dropout = 512, K, 500, "LSTM", 2, 0.5
