"""Prints the kernel timeline of one decode-graph replay (python tools/trace_decode.py [workload]).  Needs the debug build of the
library (make -C echr_b200/csrc clean && make -C echr_b200/csrc DEBUG_KNOBS=1): the shipped library ignores ECHR_TRACE."""
import os, sys
os.environ["ECHR_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from echr_b200 import synth
from echr_b200.caption_generator import CaptionGenerator
from echr_b200.config import echr_opt
from echr_b200.engine import BatchLayout

wl = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
prec = sys.argv[2] if len(sys.argv) > 2 else "bf16w"
torch.set_grad_enabled(False)
opt = echr_opt()
sd = synth.round_state_dict_bf16(synth.init_state_dict(opt, 0))
m = CaptionGenerator(opt, precision=prec); m.load_state_dict(sd); m = m.cuda()
vids, w = synth.make_workload(wl, n_videos=int(sys.argv[3]) if len(sys.argv) > 3 else None)
lay = BatchLayout([v.c3d_feats.shape[0] for v in vids], [v.ind_select_list for v in vids], [v.soi_select_list for v in vids])
c3d = torch.cat([v.c3d_feats for v in vids]).cuda(); tap = torch.cat([v.tap_feats for v in vids]).cuda()
lda = torch.stack([v.lda_feats for v in vids]).cuda()
eng = m.engine
for _ in range(3):
    eng.set_inputs(lay, c3d, tap, lda); eng.encode(); eng.decode_greedy()
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
ev[0].record(); eng.set_inputs(lay, c3d, tap, lda); ev[1].record(); eng.encode(); ev[2].record(); eng.decode_greedy(); ev[3].record()
torch.cuda.synchronize()
print("set_inputs %.3f ms, encode %.3f ms, decode %.3f ms" % (ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3])))
# the same three phases with the GPU otherwise idle, five times each (what a lane of the bench adds up to per batch)
for name, fn in (("set_inputs", lambda: eng.set_inputs(lay, c3d, tap, lda)), ("encode", eng.encode), ("decode_greedy", eng.decode_greedy)):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(5):
        fn()
    b.record(); torch.cuda.synchronize()
    print("%-14s %.3f ms per call (5 back to back)" % (name, a.elapsed_time(b) / 5))
print("time_step_kernels (each kernel alone, back to back, us):", {k: round(v * 1e3, 1) for k, v in eng.time_step_kernels(beam=1, iters=20).items()})
eng.decode_greedy(); torch.cuda.synchronize()
tr = eng.trace()
print("kernels:", len(tr))
if not tr:
    print("(no kernel trace: this is not the debug build)")
    sys.exit(0)
for i, (k, s, e) in enumerate(tr):
    if i < 14 or i >= len(tr) - 6:
        print("%4d %-11s start %9.2f  end %9.2f  dur %7.2f us" % (i, k, s, e, e - s))
import collections
dur = collections.defaultdict(list)
for k, s, e in tr:
    dur[k].append(e - s)
for k, v in dur.items():
    print("%-11s n=%3d  mean dur %.2f us" % (k, len(v), sum(v) / len(v)))
# step period: gates start to next gates start
g = [s for k, s, e in tr if k == "lstm_gates"]
print("mean gates-to-gates period: %.2f us" % ((g[-1] - g[0]) / (len(g) - 1)))
# gaps on the main chain
by = [(k, s, e) for k, s, e in tr]
main = [x for x in by if x[0] in ("lstm_gates", "logit", "finalize")]
main.sort(key=lambda x: x[1])
gaps = collections.defaultdict(list)
for a, b in zip(main, main[1:]):
    gaps[a[0] + "->" + b[0]].append(b[1] - a[2])
for k, v in gaps.items():
    print("gap %-24s mean %.2f us" % (k, sum(v) / len(v)))
side = [x for x in by if x[0] in ("h2att", "attention")]
side.sort(key=lambda x: x[1])
for a, b in list(zip(side, side[1:]))[:4]:
    print("side gap %s->%s %.2f" % (a[0], b[0], b[1] - a[2]))

det = eng.trace_detail
for kind in ("h2att", "lstm_gates", "logit"):
    rows = [det[i] for i, (k, s, e) in enumerate(tr) if k == kind and det[i][3] > 0]
    if rows:
        import numpy as np
        a = np.array(rows, dtype=np.float64)
        c = a[:, 3]
        print("%-11s per-CTA mean (us): first stage landed %.2f, accumulator ready %.2f, CTA end %.2f  (CTAs %d)"
              % (kind, (a[:, 0] / c).mean() / 1e3, (a[:, 1] / c).mean() / 1e3, (a[:, 2] / c).mean() / 1e3, int(c.mean())))
