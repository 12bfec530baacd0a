"""TEST INFRASTRUCTURE (CPU restatement, torch fp32): the SST proposal encoder.
Only tests/, smoke() and CPU-baseline measurement legs may import this.

Restates models/sst_model.py:31-40 (SST.forward in eval mode) with the LSTM written out cell by cell
(nn.LSTM semantics: gates i,f,g,o = W_ih x + b_ih + W_hh h + b_hh; c' = f*c + i*g; h' = o*tanh(c'); layer 2 reads
layer 1's h of the same step; zero initial state) instead of calling nn.LSTM.
Parity pinned: tests/golden/sst.npz holds the outputs of the reference's own SST module on the seeded weights
and inputs of tests/sst_cases.py (generator: tests/golden/make_sst_golden.py).
"""
import torch


def sst_forward(sd, features):
    """sd: tap_model state_dict (rnn.weight_ih_l0 ... scores.bias), features [T,500] fp32.
    Returns (rnn_output [T,512], sigmoid scores [T,K])."""
    x = features.to(torch.float32)
    T = x.shape[0]
    layers = 0
    while "rnn.weight_ih_l%d" % layers in sd:
        layers += 1
    seq = x
    for l in range(layers):
        w_ih, w_hh = sd["rnn.weight_ih_l%d" % l].float(), sd["rnn.weight_hh_l%d" % l].float()
        b = sd["rnn.bias_ih_l%d" % l].float() + sd["rnn.bias_hh_l%d" % l].float()
        H = w_hh.shape[1]
        h = torch.zeros(H)
        c = torch.zeros(H)
        pre = seq @ w_ih.t() + b  # input projection of every clip
        out = []
        for t in range(T):
            g = pre[t] + w_hh @ h
            i, f, gg, o = g[:H], g[H:2 * H], g[2 * H:3 * H], g[3 * H:]
            c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(gg)
            h = torch.sigmoid(o) * torch.tanh(c)
            out.append(h)
        seq = torch.stack(out) if out else torch.zeros(0, H)
    scores = torch.sigmoid(seq @ sd["scores.weight"].float().t() + sd["scores.bias"].float())
    return seq, scores
