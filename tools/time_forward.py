"""Wall-clock of the drop-in call (CaptionGenerator.forward, one video per call) as eval_utils.py:129 makes it."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from echr_b200 import synth
from echr_b200.caption_generator import CaptionGenerator
from echr_b200.config import echr_opt

torch.set_grad_enabled(False)
opt = echr_opt()
sd = synth.init_state_dict(opt, 0)
m = CaptionGenerator(opt, precision=sys.argv[1] if len(sys.argv) > 1 else "fp32"); m.load_state_dict(sd); m = m.cuda()
# an eval-loop-like stream: every video has its own clip count and number of events
vids = [synth.make_video(i, T=60 + 11 * (i % 29), n_events=1 + (i * 7) % 9, style="gt") for i in range(64)]
args = [v.args("cuda:0") for v in vids]
for a in args[:8]:
    m(*a, mode="eval")
torch.cuda.synchronize()
t0 = time.perf_counter()
n = 0
for rep in range(3):
    for a in args:
        seq, lp = m(*a, mode="eval")
        s = lp.sum(1).cpu()  # eval_utils.py:134
        n += 1
torch.cuda.synchronize()
dt = time.perf_counter() - t0
print("forward(mode='eval'), one video per call: %.3f ms per call (%d calls), %.0f events/s" % (dt / n * 1e3, n, sum(len(v.ind_select_list) for v in vids) * 3 / dt))

import cProfile, pstats, io
pr = cProfile.Profile()
pr.enable()
for a in args:
    seq, lp = m(*a, mode="eval")
    s = lp.sum(1).cpu()
pr.disable()
st = io.StringIO()
pstats.Stats(pr, stream=st).sort_stats("cumulative").print_stats(25)
print(st.getvalue())
