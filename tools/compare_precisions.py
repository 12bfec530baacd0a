"""Token agreement of a faster precision with the fp32-grade one on the GPU, over the videos of a bench workload.
For every event whose greedy tokens diverge, the top-2 gap of the fp32-grade path at the first divergent step
is reported: BASELINE's rule excuses a divergence only where that gap is below 1e-3.
usage: python tools/compare_precisions.py [precision=f16] [workload=cfg2] [n_videos]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from echr_b200 import synth
from echr_b200.caption_generator import CaptionGenerator
from echr_b200.config import echr_opt

prec = sys.argv[1] if len(sys.argv) > 1 else "f16"
wl = sys.argv[2] if len(sys.argv) > 2 else "cfg2"
nv = int(sys.argv[3]) if len(sys.argv) > 3 else None
torch.set_grad_enabled(False)
opt = echr_opt()
sd = synth.round_state_dict_bf16(synth.init_state_dict(opt, 0))
ref = CaptionGenerator(opt, precision="fp32"); ref.load_state_dict(sd); ref = ref.cuda()
fast = CaptionGenerator(opt, precision=prec); fast.load_state_dict(sd); fast = fast.cuda()
vids = synth.make_workload(wl, n_videos=nv)[0] if nv else synth.make_workload(wl)[0]
events = diverged = unexcused = 0
worst_logp = worst_tf = 0.0
gaps = []
for v in vids:
    a = (v.tap_feats.cuda(), v.c3d_feats.cuda(), v.lda_feats.cuda(), [], list(v.ind_select_list), [list(s) for s in v.soi_select_list])
    s0, l0 = ref(*a, mode="eval")
    s1, l1 = fast(*a, mode="eval")
    lab = torch.cat([torch.zeros(s0.shape[0], 1, dtype=torch.long, device=s0.device), s0], 1)
    tf0 = ref(a[0], a[1], a[2], lab, a[4], a[5], mode="train")
    tf1 = fast(a[0], a[1], a[2], lab, a[4], a[5], mode="train")
    n = min(tf0.shape[0], tf1.shape[0])
    used = (lab[:n, :tf0.shape[1]] > 0) | (torch.arange(tf0.shape[1], device=lab.device)[None, :] == 0)
    d = (tf0[:n] - tf1[:n]).abs().amax(2)
    worst_tf = max(worst_tf, float(d[used[:, :d.shape[1]]].max()))
    top2 = tf0.topk(2, dim=2)[0]
    s0c, s1c = s0.cpu().numpy(), s1.cpu().numpy()
    for r in range(s0c.shape[0]):
        events += 1
        diff = np.nonzero(s0c[r] != s1c[r])[0]
        if diff.size == 0:
            worst_logp = max(worst_logp, float((l0[r] - l1[r]).abs().max()))
            continue
        t = int(diff[0])
        gap = float(top2[r, t, 0] - top2[r, t, 1])
        gaps.append(gap)
        diverged += 1
        unexcused += gap >= 1e-3
print("precision %s vs fp32 on %s: %d events, %d diverge (%d with top-2 gap >= 1e-3: %s), max |dlogp| greedy %.2e, teacher-forced %.2e"
      % (prec, wl, events, diverged, unexcused, ["%.1e" % g for g in sorted(gaps)], worst_logp, worst_tf))
sys.exit(1 if unexcused else 0)
