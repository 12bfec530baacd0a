"""Experiment: throughput with 1, 2, 3 batches in flight (one engine + stream per slot), device- and host-resident inputs.
usage: python tools/two_in_flight.py [precision] [workload] [n_videos]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from echr_b200 import synth
from echr_b200.config import echr_opt
from echr_b200.engine import BatchLayout, Engine

prec = sys.argv[1] if len(sys.argv) > 1 else "f16"
wl = sys.argv[2] if len(sys.argv) > 2 else "cfg2"
nv = int(sys.argv[3]) if len(sys.argv) > 3 else None
torch.set_grad_enabled(False)
opt = echr_opt()
sd = synth.round_state_dict_bf16(synth.init_state_dict(opt, 0))
vids = synth.make_workload(wl, n_videos=nv)[0] if nv else synth.make_workload(wl)[0]
lay = BatchLayout([v.c3d_feats.shape[0] for v in vids], [v.ind_select_list for v in vids], [v.soi_select_list for v in vids])
h = [torch.cat([v.c3d_feats for v in vids]).pin_memory(), torch.cat([v.tap_feats for v in vids]).pin_memory(),
     torch.stack([v.lda_feats for v in vids]).pin_memory()]
d = [t.cuda() for t in h]
S = opt.CG_seq_length
K = 40
for slots in (1, 2, 3):
    engs = [Engine(sd, opt.CG_vocab_size, S, precision=prec) for _ in range(slots)]
    streams = [torch.cuda.Stream() for _ in range(slots)]
    outs_d = [(torch.zeros(lay.n_events, S, dtype=torch.int32, device="cuda"), torch.zeros(lay.n_events, S, device="cuda"),
               torch.zeros(lay.n_events, device="cuda")) for _ in range(slots)]
    outs_h = [(torch.zeros(lay.n_events, S, dtype=torch.int32).pin_memory(), torch.zeros(lay.n_events, S).pin_memory(),
               torch.zeros(lay.n_events).pin_memory()) for _ in range(slots)]
    for host in (0, 1):
        def step(i):
            s = i % slots
            with torch.cuda.stream(streams[s]):
                if host:
                    engs[s].caption_host(lay, h[0], h[1], h[2], beam=1, out=outs_h[s], sync=False)
                else:
                    engs[s].caption_device(lay, d[0], d[1], d[2], beam=1, out=outs_d[s])
        for i in range(4 * slots):
            step(i)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in streams:
            s.wait_event(e0)
        for i in range(K):
            step(i)
        for s in streams:
            e = torch.cuda.Event(); e.record(s); torch.cuda.current_stream().wait_event(e)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / K
        print("%s %s slots=%d %s: %.4f ms per batch, %.0f events/s (wall %.4f ms)"
              % (prec, wl, slots, "host" if host else "device", ms, lay.n_events / ms * 1e3, (time.perf_counter() - t0) * 1e3 / K))
    ref = outs_d[0][0].clone()
    assert all(torch.equal(o[0], ref) for o in outs_d) and all(torch.equal(o[0].cuda(), ref) for o in outs_h)
    del engs
