"""Measurement of the SST proposal-encoder row: V videos x T clips through echr_sst_forward
(CUDA events around 10 calls, inputs resident in HBM) beside the same math on the host cores
(torch nn.LSTM + Linear, all threads, one video per call like eval_utils.py:55).  Prints one JSON line.
usage: python tools/bench_sst.py [precision] [n_videos] [T]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch
import sst_cases as C
from echr_b200.sst_model import SST

prec = sys.argv[1] if len(sys.argv) > 1 else "f16"
V = int(sys.argv[2]) if len(sys.argv) > 2 else 64
T = int(sys.argv[3]) if len(sys.argv) > 3 else 100
torch.set_grad_enabled(False)
sd = {k: v.bfloat16().float() for k, v in C.state_dict().items()}
m = SST(C.Opt(), precision=prec); m.load_state_dict(sd); m = m.cuda(); m.eval()
rng = np.random.RandomState(0)
feats = [torch.from_numpy(rng.randn(T, 500).astype(np.float32)).cuda() for _ in range(V)]
for _ in range(3):
    m.forward_batch(feats)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    out = m.forward_batch(feats)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
# host cores: stock torch modules holding the same weights
cpu = torch.nn.LSTM(500, 512, 2, batch_first=True); lin = torch.nn.Linear(512, C.K)
cpu.load_state_dict({k[4:]: v for k, v in sd.items() if k.startswith("rnn.")}); lin.load_state_dict({"weight": sd["scores.weight"], "bias": sd["scores.bias"]})
torch.set_num_threads(os.cpu_count() or 1)
hf = [f.cpu() for f in feats[:8]]
t0 = time.perf_counter()
for f in hf:
    o, _ = cpu(f.unsqueeze(0)); s = torch.sigmoid(lin(o[0]))
cpu_ms = (time.perf_counter() - t0) / len(hf) * V * 1e3
err = float((out[0][1].cpu() - torch.sigmoid(lin(cpu(hf[0].unsqueeze(0))[0][0]))).abs().max())
flops = 2.0 * V * T * 2048 * (512 + 512 + 512 + 512) + 2.0 * V * T * 512 * C.K
print(json.dumps({"row": "N1 SST proposal encoder", "precision": prec, "videos": V, "T": T, "K": C.K,
                  "gpu_ms_per_batch": ms, "clips_per_s": V * T / ms * 1e3, "algorithmic_tflops": flops / (ms * 1e-3) / 1e12,
                  "cpu_ms_per_batch": cpu_ms, "cpu": {"kind": "torch nn.LSTM + Linear, one video per call", "cores": os.cpu_count(), "sample": "8 videos scaled to %d" % V},
                  "max_abs_diff_vs_cpu_scores": err,
                  "note": "2*T sequential recurrent GEMM launches of %d rows each: launch/latency-bound, not tensor-bound" % V}))
