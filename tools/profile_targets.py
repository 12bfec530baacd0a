"""A short program that launches every kernel family of the library a few times, for `ncu --set full` (tools/gpu_round.sh
ncufull): decode with 3-token captions so that a whole decode is ~25 launches.

  1. cfg5 shape, one batch of 512 videos (~2040 rows), greedy, f16: persistent gate GEMM, persistent logit GEMM with the
     attention warps, h2att tiles, finalize, encode kernels;
  2. cfg2 (64 videos, 251 rows), greedy, f16: the latency-regime GEMM tiles and the stand-alone 16-bit attention;
  3. cfg4 shape, 4 videos x 200 proposals x T=400, greedy: fused pair kernel (1250 tiles), relation attention, long windows;
  4. cfg3 shape, 8 videos x 100 proposals, beam 3: finalize_beam, top-4 logit epilogue;
  5. proposal selection (top-N, NMS) and the SST encoder on 8 videos x 16 clips.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
from echr_b200 import proposals as P  # noqa: E402
from echr_b200 import synth  # noqa: E402
from echr_b200.caption_generator import CaptionGenerator  # noqa: E402
from echr_b200.config import echr_opt  # noqa: E402
from echr_b200.sst_model import SST  # noqa: E402

torch.set_grad_enabled(False)
DEV = "cuda:0"
opt = echr_opt(CG_seq_length=3)
sd = synth.round_state_dict_bf16(synth.init_state_dict(opt, seed=0))
m = CaptionGenerator(opt, precision="f16")
m.load_state_dict(sd, strict=True)
m = m.to(DEV)
eng = m.engine


def run(vids, beam, what):
    lay, c3d, tap, lda = helpers.batch_of(vids, DEV)
    tok = eng.caption_device(lay, c3d, tap, lda, beam=beam)[0]
    torch.cuda.synchronize()
    print(what, "rows", lay.n_events * beam, "tokens", tuple(tok.shape))


if "--step-only" in sys.argv:  # only the default bench's batch shape (1024 videos): the step kernels' traffic for bench.py
    run(synth.make_workload("cfg5", n_videos=1024, bf16_features=True)[0], 1, "cfg5 batch of 1024 videos")
    sys.exit(0)
run(synth.make_workload("cfg5", n_videos=512, bf16_features=True)[0], 1, "cfg5 batch")
run(synth.make_workload("cfg2", bf16_features=True)[0], 1, "cfg2")
run(synth.make_workload("cfg4", n_videos=4, bf16_features=True)[0], 1, "cfg4 x4")
run(synth.make_workload("cfg3", n_videos=8, bf16_features=True)[0], 3, "cfg3 x8 beam 3")


class SOpt(object):
    hidden_dim, K, video_dim, tap_rnn_type, rnn_num_layers, rnn_dropout = 512, 64, 500, "LSTM", 2, 0.5


rng = np.random.RandomState(1)
sst = SST(SOpt(), precision="f16")
ssd = {k: torch.from_numpy(((rng.rand(*t.shape) * 2 - 1) / np.sqrt(512.0)).astype(np.float32)) for k, t in sst.state_dict().items()}
sst.load_state_dict(ssd, strict=True)
sst = sst.to(DEV)
sst.eval()
feats = [torch.randn(16, 500, device=DEV) for _ in range(8)]
outs = sst.forward_batch(feats)
torch.cuda.synchronize()
T, K = 100, 64
pred = torch.rand(T, K, device=DEV)
mask = (np.arange(T)[:, None] >= np.arange(K)[None, :]).astype(np.float32)
P.gettop1000(pred, mask, [], 100.0, lambda s, e, n, d: (s * d / n, e * d / n), topN=100)
P.gettop1000_nms(pred, mask, [], 100.0, lambda s, e, n, d: (s * d / n, e * d / n), topN=50, overlap=0.7)
torch.cuda.synchronize()
print("done")
