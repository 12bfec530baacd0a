"""Generates tests/golden/*.npz by running the UNMODIFIED reference (through oracle/ref_shim.py) in
the authoring container.  The fixtures pin both the oracle (tests/test_oracle.py) and the CUDA path
(tests/test_golden_gpu.py) on machines where /root/reference does not exist.

    python tests/golden/make_golden.py

Inputs are not stored: they are regenerated from seeds by echr_b200/synth.py (torch CPU generator).
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from echr_b200 import synth  # noqa: E402
from echr_b200.config import echr_opt  # noqa: E402
from oracle import ref_shim  # noqa: E402

# name -> (video seed, T, n_events, style)
CASES = {
    "cfg1": (0, 100, 3, "gt"),
    "gt5": (1, 100, None, "gt"),
    "sst12": (2, 100, 12, "sst"),
    "long30": (3, 400, 30, "sst"),
}
TF_STEPS = (0, 7, 29)
TF_EVENTS = 2


def main():
    torch.set_grad_enabled(False)
    torch.manual_seed(0)
    opt = echr_opt()
    for wname, rounder in (("fp32", None), ("bf16w", synth.round_state_dict_bf16)):
        sd = synth.init_state_dict(opt, seed=0)
        if rounder:
            sd = rounder(sd)
        model = ref_shim.build_reference(opt, sd)
        for cname, (seed, T, n, style) in CASES.items():
            v = synth.make_video(seed, T=T, n_events=n, style=style, bf16_features=(wname == "bf16w"))
            args = v.args()
            seq, logp = model(*args, mode="eval")
            event = model.get_event_context(args[0], args[1], args[2], args[4], args[5])
            labels = torch.cat([torch.zeros(seq.shape[0], 1, dtype=torch.long), seq], 1)
            tf = model(args[0], args[1], args[2], labels, args[4], args[5], mode="train")
            assert tf.shape[1] == seq.shape[1], (tf.shape, seq.shape)
            out = os.path.join(ROOT, "tests", "golden", "%s_%s.npz" % (cname, wname))
            np.savez_compressed(
                out, seq=seq.numpy().astype(np.int32), logp=logp.numpy(), event=event.numpy(),
                tf_steps=np.array(TF_STEPS), tf_logprobs=tf[:TF_EVENTS][:, list(TF_STEPS)].numpy(),
                tf_top2=tf.topk(2, dim=2)[0].numpy(), meta=np.array([seed, T, v.n_events]))
            print(out, seq.shape, os.path.getsize(out))


if __name__ == "__main__":
    main()
