"""Generates tests/golden/sst.npz: outputs of the UNMODIFIED reference module models/sst_model.py:SST (eval mode)
on the seeded weights / inputs of tests/sst_cases.py.  Run where /root/reference exists:

    python tests/golden/make_sst_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, "/root/reference")
from models.sst_model import SST  # noqa: E402  (the reference)
import sst_cases as C  # noqa: E402


def main():
    torch.set_grad_enabled(False)
    torch.set_num_threads(1)
    m = SST(C.Opt())
    m.load_state_dict(C.state_dict(), strict=True)
    m.eval()  # the reference's own eval(): rnn.dropout = 0
    out = {}
    for c in range(len(C.CASE_LENGTHS)):
        rnn_output, scores = m(C.features(c))
        out["tap%d" % c] = rnn_output.numpy()
        out["pred%d" % c] = scores.numpy()
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "sst.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
