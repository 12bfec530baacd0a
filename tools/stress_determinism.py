"""Race hunt: replays the same batch many times and compares tokens / log-probs bitwise with the first replay.
usage: python tools/stress_determinism.py [iters] [precision = f16] ; the environment switches (ECHR_PDL_MASK, ECHR_KSPLIT, ...) apply in a DEBUG_KNOBS=1 build only."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from echr_b200 import synth
from echr_b200.caption_generator import CaptionGenerator
from echr_b200.config import echr_opt
from echr_b200.engine import BatchLayout

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 200
precision = sys.argv[2] if len(sys.argv) > 2 else "f16"
torch.set_grad_enabled(False)
opt = echr_opt()
sd = synth.round_state_dict_bf16(synth.init_state_dict(opt, 0))
m = CaptionGenerator(opt, precision=precision); m.load_state_dict(sd); m = m.cuda()
eng = m.engine
cases = {
    "5ev": [synth.make_video(7, n_events=5, style="gt")],
    "18ev": [synth.make_video(20 + i, n_events=n, style=st) for i, (n, st) in enumerate([(3, "gt"), (1, "gt"), (7, "sst"), (5, "gt"), (2, "sst")])],
    "cfg2": synth.make_workload("cfg2")[0],
    "300ev": synth.make_workload("cfg2", n_videos=76)[0],
    "1200ev": synth.make_workload("cfg2", n_videos=300)[0],
}
print("precision", precision)
bad_total = 0
for name, vids in cases.items():
    lay = BatchLayout([v.c3d_feats.shape[0] for v in vids], [v.ind_select_list for v in vids], [v.soi_select_list for v in vids])
    c3d = torch.cat([v.c3d_feats for v in vids]).cuda(); tap = torch.cat([v.tap_feats for v in vids]).cuda()
    lda = torch.stack([v.lda_feats for v in vids]).cuda()
    for beam in (1, 3):
        ref = None
        bad = 0
        worst = 0.0
        for it in range(iters if lay.n_events < 1000 else max(20, iters // 6)):
            tok, lp, sc = eng.caption_device(lay, c3d, tap, lda, beam=beam)
            tok, lp = tok.clone(), lp.clone()
            if ref is None:
                ref = (tok, lp)
            elif not (torch.equal(tok, ref[0]) and torch.equal(lp, ref[1])):
                bad += 1
                worst = max(worst, float((lp - ref[1]).abs().max()))
        print("%-6s beam %d rows %5d: %d / %d replays differ from the first (max |dlogp| %.3e)" % (name, beam, lay.n_events * beam, bad, (iters if lay.n_events < 1000 else max(20, iters // 6)) - 1, worst), flush=True)
        bad_total += bad
print("TOTAL mismatching replays:", bad_total)
sys.exit(1 if bad_total else 0)
