"""SST proposal encoder on the B200: host-side mirror of the reference's
models/sst_model.py — same class name, constructor options, parameter names (a reference `tap_model`
checkpoint loads with strict=True) and `forward(features) -> (rnn_output, outputs)`.  The math runs in
csrc/sst.cu (tcgen05 GEMMs with the LSTM epilogue, time-major recurrence) behind echr_sst_forward; the
nn.LSTM / nn.Linear members only hold the parameters.  `forward_batch` runs many videos of different length
in one call, which is what the reference's per-video loop (eval_utils.py:55) cannot do.
Inference only: training (dropout between the LSTM layers, backward) is out of scope, and there is no CPU path.
"""
import ctypes

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from ._lib import EchrError, check
from .engine import PRECISIONS

SST_WEIGHT_FIELDS = (
    ("w_ih0", "rnn.weight_ih_l0"), ("w_hh0", "rnn.weight_hh_l0"), ("b_ih0", "rnn.bias_ih_l0"), ("b_hh0", "rnn.bias_hh_l0"),
    ("w_ih1", "rnn.weight_ih_l1"), ("w_hh1", "rnn.weight_hh_l1"), ("b_ih1", "rnn.bias_ih_l1"), ("b_hh1", "rnn.bias_hh_l1"),
    ("scores_w", "scores.weight"), ("scores_b", "scores.bias"),
)


class SstWeights(ctypes.Structure):
    _fields_ = [(f, ctypes.c_void_p) for f, _ in SST_WEIGHT_FIELDS]


class SST(nn.Module):
    """models/sst_model.py:5-40."""

    def __init__(self, opt, precision="fp32"):
        super(SST, self).__init__()
        if precision not in PRECISIONS:
            raise ValueError("precision must be one of %s" % sorted(PRECISIONS))
        if opt.video_dim != 500 or opt.hidden_dim != 512 or opt.rnn_num_layers != 2 or getattr(opt, "tap_rnn_type", "LSTM") != "LSTM":
            raise EchrError("the B200 SST encoder is built for the ECHR recipe: LSTM, video_dim 500, hidden_dim 512, 2 layers")
        self.scores = torch.nn.Linear(opt.hidden_dim, opt.K)
        self.video_dim = opt.video_dim
        self.rnn_type = opt.tap_rnn_type
        self.rnn_num_layers = opt.rnn_num_layers
        self.rnn_dropout = opt.rnn_dropout
        self.K = opt.K
        self.data_for_test = []
        self.rnn = nn.LSTM(opt.video_dim, opt.hidden_dim, opt.rnn_num_layers, batch_first=True, dropout=opt.rnn_dropout)
        self.precision = precision
        self._handle = None
        self._packed_for = None

    # the reference overrides these to toggle the inter-layer dropout only (sst_model.py:25-29)
    def eval(self):
        self.rnn.dropout = 0

    def train(self, mode=True):
        self.rnn.dropout = self.rnn_dropout if mode else 0

    # ------------------------------------------------------------------ packed weights
    def _close(self):
        if self._handle is not None:
            _lib.load().echr_sst_destroy(self._handle)
            self._handle = None

    def __del__(self):
        try:
            self._close()
        except Exception:
            pass

    def _engine(self, device):
        version = tuple(p._version for p in self.parameters()) + (str(device),)
        if self._handle is not None and self._packed_for == version:
            return self._handle
        self._close()
        if device.type != "cuda":
            raise EchrError("echr_b200 runs on a B200 (sm_100a) only; there is no CPU path")
        lib = _lib.load()
        sd = self.state_dict()
        ws, keep = SstWeights(), []
        for field, key in SST_WEIGHT_FIELDS:
            t = sd[key].detach().to(device=device, dtype=torch.float32).contiguous()
            keep.append(t)
            setattr(ws, field, t.data_ptr())
        a, w, _ = PRECISIONS[self.precision]
        h = ctypes.c_void_p()
        with torch.cuda.device(device):
            check(lib.echr_sst_create(int(self.K), a, w, ctypes.byref(ws), torch.cuda.current_stream().cuda_stream,
                                      ctypes.byref(h)), "echr_sst_create")
        self._handle, self._packed_for = h, version
        return h

    # ------------------------------------------------------------------ forward
    def forward_batch(self, features_list):
        """features_list: fp32 CUDA tensors [T_v, 500].  Returns ([T_v,512] rnn_output, [T_v,K] scores) per video,
        all videos through one library call."""
        if self.rnn.dropout:
            raise NotImplementedError("training-mode dropout is out of scope: call eval() (sst_model.py:25-26)")
        if hasattr(self, "reduce_dim_layer"):
            features_list = [self.reduce_dim_layer(f) for f in features_list]
        dev = features_list[0].device
        h = self._engine(dev)
        lens = [int(f.shape[0]) for f in features_list]
        off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
        c3d = torch.cat([f.to(device=dev, dtype=torch.float32) for f in features_list]).contiguous()
        if c3d.shape[1] != 500:
            raise ValueError("features must be [T, 500], got %s" % (tuple(c3d.shape),))
        n = int(off[-1])
        tap = torch.zeros(n, 512, dtype=torch.float32, device=dev)
        pred = torch.zeros(n, self.K, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            check(_lib.load().echr_sst_forward(h, c3d.data_ptr(), off.ctypes.data, len(lens), tap.data_ptr(), pred.data_ptr(),
                                               torch.cuda.current_stream().cuda_stream), "echr_sst_forward")
        return [(tap[off[i]:off[i + 1]], pred[off[i]:off[i + 1]]) for i in range(len(lens))]

    def forward(self, features):
        """sst_model.py:31-40: features [T,500] -> (rnn_output [T,512], outputs [T,K])."""
        return self.forward_batch([features])[0]
