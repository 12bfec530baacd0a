#!/usr/bin/env python
"""Benchmark of the ECHR caption-generation hot path (BASELINE.json metric: events captioned / s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg5] [--beam B] [--precision f16]
                    [--in-flight 4] [--batch-videos 1024]
    python bench.py --impl reference ...      # the reference's own CPU path on the host cores (oracle/_ref)

Default workload = BASELINE.json configs[4] ("cfg5"): 4096 synthetic videos (~16k events) in total, block-sharded
by video over the N ranks (strong scaling; videos are independent, so there is no data-path collective), every
rank feeding its shard batch after batch through a CaptionPipeline, one all_gather of the decoded token ids per
batch.  One "step" = the whole workload captioned once.  `value` is greedy decode with the inputs resident in
HBM; `e2e` the same through the host-buffer call (BatchLayout construction, H2D of every input and D2H of the
results inside the timed region); `beam3` the same two numbers for beam = 3.  cfg2/cfg3/cfg4 (`--workload`) are the
other BASELINE configurations.  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from echr_b200 import synth  # noqa: E402
from echr_b200.config import echr_opt  # noqa: E402

METRIC = "events_captioned_per_sec"
UNIT = "events/s"
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "r02_traffic.json")
# videos of the workload in the bounded CPU sample (about 10-30 s of CPU work for the default step counts)
CPU_SAMPLE = {"cfg1": 1, "cfg2": 64, "cfg3": 1, "cfg4": 1, "cfg5": 64}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg5", choices=sorted(synth.WORKLOADS))
    ap.add_argument("--videos", type=int, default=None, help="override the workload's TOTAL video count")
    ap.add_argument("--beam", type=int, default=None, help="beam of the headline value (default: the workload's)")
    ap.add_argument("--precision", default="f16", choices=["fp32", "bf16w", "f16"])
    ap.add_argument("--cpu-sample", type=int, default=None, help="videos in the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-second-beam", action="store_true", help="skip the beam=3 (greedy for cfg3) companion measurement")
    ap.add_argument("--in-flight", type=int, default=4,
                    help="batches kept in flight per GPU (CaptionPipeline lanes); 1 = one blocking batch at a time")
    ap.add_argument("--batch-videos", type=int, default=1024, help="videos per batch a rank submits")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        return json.load(open(path)), "measured (MEASURED_PEAKS.json)"
    # B200_PROFILING.md fallback figures
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback (B200_PROFILING.md)"


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def stop(self, windows=()):
        """windows: (t0, t1) host-clock intervals of the timed regions; the median is over samples inside them."""
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        inside = [r for t, r in self.rows if any(a <= t <= b + 0.1 for a, b in windows)] or [r for _, r in self.rows]
        sm = sorted(int(r[0]) for r in inside if r and r[0].isdigit())
        mx = [int(r[1]) for _, r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i] == "Active" for _, r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "samples_total": len(self.rows)}


def nvml_state(index):
    """One NVML reading of this rank's GPU (diagnosis line on stderr under torchrun)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        return "sm %d MHz, mem %d MHz, %.0f W, event reasons 0x%x" % (
            pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_MEM),
            pynvml.nvmlDeviceGetPowerUsage(h) / 1e3, pynvml.nvmlDeviceGetCurrentClocksEventReasons(h))
    except Exception as e:  # noqa: BLE001  (diagnosis only)
        return "nvml unavailable (%s)" % type(e).__name__


# ---------------------------------------------------------------------------------------------- CPU arm
def cpu_reference_run(vids, sd, beam, reps, warm):
    """The reference's own work shape on the host cores, all threads, one video per call (eval_utils.py:38-39,129).
    Runs the reference's modules themselves when they are reachable (oracle/_ref staged by oracle/make_ref.py, or
    /root/reference in the authoring container): kind "reference"; otherwise the oracle port: kind "port".
    Beam (the reference has none): the oracle's beam loop driving the reference's own
    get_logprobs_state.  Returns (events/s, seconds per pass, kind)."""
    from oracle import echr_oracle as O
    from oracle import ref_shim
    torch.set_num_threads(os.cpu_count() or 1)
    torch.set_grad_enabled(False)
    n_events = sum(v.n_events for v in vids)
    ref = ref_shim.build_reference(echr_opt(), sd) if ref_shim.available() else None

    def one_pass():
        for v in vids:
            if ref is None:
                if beam > 1:
                    O.caption_video_beam(sd, v.tap_feats, v.c3d_feats, v.lda_feats, v.ind_select_list, v.soi_select_list, beam=beam)
                else:
                    O.caption_video(sd, v.tap_feats, v.c3d_feats, v.lda_feats, v.ind_select_list, v.soi_select_list)
            elif beam > 1:
                a = v.args()
                video = ref.get_video_context(a[0], a[1], a[2], a[4], a[5])
                event = ref.get_event_context(a[0], a[1], a[2], a[4], a[5])
                clip, mask = ref.get_clip_context(a[0], a[1], a[2], a[4], a[5])
                lm = ref.lm_model
                O.beam_search(sd, video, event, clip, mask, beam=beam,
                              step_fn=lambda it, ev, cl, mk, st: lm.get_logprobs_state(it, video, ev, cl, mk, st))
            else:
                ref(*v.args(), mode="eval")

    for _ in range(warm):
        one_pass()
    times = []
    for _ in range(reps):
        t0 = time.perf_counter()
        one_pass()
        times.append(time.perf_counter() - t0)
    times.sort()
    med = times[len(times) // 2]
    return n_events / med, med, ("reference" if ref is not None else "port")


def workload_plan(args):
    """Everything both arms agree on: the videos' event counts (features are not built here) and `config`."""
    w = dict(synth.WORKLOADS[args.workload])
    if args.videos is not None:
        w["n_videos"] = args.videos
    if args.beam is not None:
        w["beam"] = args.beam
    events = [synth.video_events(i, T=w["T"], n_events=w["n_events"], style=w["style"]) for i in range(w["n_videos"])]
    counts = [len(e[0]) for e in events]
    config = {"workload": "%s: %d videos x T=%d clips, %d events (%s proposals), beam=%d, vocab 5747, max_len 30"
                          % (args.workload, w["n_videos"], w["T"], sum(counts), w["style"], w["beam"]),
              "videos": w["n_videos"], "events": int(sum(counts)), "beam": w["beam"],
              "parallelism": "videos block-sharded over the ranks (balanced by events); one all_gather of token ids per batch",
              "l2": "a 256 MB write (> 126 MB L2) at the head of every step, inside the timed region; at N=1 the "
                    "inputs of a step (415 MB for cfg5) also exceed L2"}
    return w, events, counts, config


def main_reference(args, rank, world):
    if rank != 0:
        return 0
    w, events, counts, config = workload_plan(args)
    sd = synth.init_state_dict(echr_opt(), seed=0)
    sample = min(args.cpu_sample or CPU_SAMPLE[args.workload], w["n_videos"])
    sub = [synth.make_video(i, T=w["T"], n_events=w["n_events"], style=w["style"]) for i in range(sample)]
    n_ev = sum(v.n_events for v in sub)
    value, sec, kind = cpu_reference_run(sub, sd, w["beam"], reps=max(1, args.steps), warm=max(0, args.warmup))
    cores = os.cpu_count() or 1
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": "each step = the first %d videos (%d events) of the workload, one video per call, "
                                       "torch CPU fp32 on %d threads; median of %d steps" % (sample, n_ev, cores, max(1, args.steps))},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    _RESULT.append(json.dumps(line))
    return 0


def main():
    args = parse()
    # stdout carries exactly one JSON line: anything a library prints there meanwhile (NCCL's INFO lines, ...) is
    # sent to stderr; the real stdout comes back for the final print
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        return _main(args)
    finally:
        # fd 1 stays pointed at stderr for the rest of the process (NCCL still prints while it tears down at exit);
        # the line goes straight to the real stdout
        sys.stdout.flush()
        if _RESULT:
            os.write(real_stdout, (_RESULT[0] + "\n").encode())
        os.close(real_stdout)


_RESULT = []


class Batch(object):
    """One batch of a rank's shard: python-side proposal lists (the reference's calling convention), pinned host
    features, resident device copies."""

    def __init__(self, vids, dev):
        self.clip_counts = [v.c3d_feats.shape[0] for v in vids]
        self.ind_lists = [v.ind_select_list for v in vids]
        self.soi_lists = [v.soi_select_list for v in vids]
        self.h = (torch.cat([v.c3d_feats for v in vids]).contiguous().pin_memory(),
                  torch.cat([v.tap_feats for v in vids]).contiguous().pin_memory(),
                  torch.stack([v.lda_feats for v in vids]).contiguous().pin_memory())
        self.d = tuple(t.to(dev) for t in self.h)
        self.n_videos = len(vids)
        self.n_events = sum(len(s) for s in self.soi_lists)
        self.windows = [(off, s, e) for off, sois in zip(np.cumsum([0] + self.clip_counts[:-1]), self.soi_lists) for s, e in sois]

    def layout(self):
        from echr_b200.engine import BatchLayout
        return BatchLayout(self.clip_counts, self.ind_lists, self.soi_lists)

    def h2d_bytes(self):
        return int(self.h[0].numel() + self.n_events * 512 + self.h[2].numel()) * 4 + int(9 * self.n_events + 2 * self.n_videos) * 4

    def unique_window_clips(self):
        """Clips of the batch that lie inside at least one proposal (the attention never reads the others)."""
        covered = np.zeros(int(sum(self.clip_counts)), dtype=bool)
        for off, s, e in self.windows:
            covered[off + s:off + e] = True
        return int(covered.sum())


def _main(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return main_reference(args, rank, world)

    import torch.distributed as dist
    from echr_b200.caption_generator import CaptionGenerator
    from echr_b200.engine import CaptionPipeline
    from echr_b200.parallel import JobPlan, gather_tokens

    torch.set_grad_enabled(False)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    opt = echr_opt()
    sd_full = synth.init_state_dict(opt, seed=0)
    sd = synth.round_state_dict_bf16(sd_full) if args.precision != "fp32" else sd_full
    model = CaptionGenerator(opt, precision=args.precision)
    model.load_state_dict(sd, strict=True)
    model = model.to(dev)
    eng = model.engine
    S = opt.CG_seq_length

    # ---- the job: every rank derives the same plan from the videos' event counts
    w, events, counts_all, config = workload_plan(args)
    job = JobPlan(counts_all, world, args.batch_videos)  # the same plan on every rank, no communication
    n_batches, plan, ev_counts = job.n_batches, job.batches, job.events
    feats16 = args.precision != "fp32"

    def build(a, b):
        return [synth.make_video(i, T=w["T"], n_events=w["n_events"], style=w["style"], bf16_features=feats16) for i in range(a, b)]

    mine = [Batch(build(a, b), dev) for a, b in plan[rank] if b > a]
    mine_idx = [bi for bi, (a, b) in enumerate(plan[rank]) if b > a]
    assert len(mine) == n_batches or world > 1, "single-rank plan has an empty batch"
    n_events = sum(bt.n_events for bt in mine)
    total_events = int(sum(counts_all))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    align = torch.zeros(1, device=dev)
    pipe = CaptionPipeline(eng, depth=max(1, args.in_flight))
    empty_tok = torch.zeros(0, S, dtype=torch.int32, device=dev)

    def gather_fn(bi, on_host):
        if world == 1:
            return None
        cnt = [ev_counts[r][bi] for r in range(world)]

        def fn(res):
            tok = res[0].to(dev, non_blocking=True) if on_host else res[0]
            return (gather_tokens(tok, counts=cnt),) + tuple(res[1:])
        return fn

    def one_step(on_host, beam, keep=None):
        """The whole shard once.  on_host: BatchLayout construction + H2D + D2H inside; else inputs resident."""
        first = True
        for slot in range(n_batches):
            if slot in mine_idx:
                bt = mine[mine_idx.index(slot)]
                lay = bt.layout() if on_host else bt.lay
                src = bt.h if on_host else bt.d
                t = pipe.submit(lay, src[0], src[1], src[2], beam=beam, before=flush.zero_ if first else None,
                                after=gather_fn(slot, on_host))
                first = False
                if keep is not None:
                    keep.append((slot, t))
            elif world > 1:  # this rank has no videos in this slot: it still takes part in the slot's gather
                cnt = [ev_counts[r][slot] for r in range(world)]
                out = gather_tokens(empty_tok, counts=cnt)
                if keep is not None:
                    keep.append((slot, out))

    for bt in mine:
        bt.lay = bt.layout()

    import math
    # every (lane, batch shape) pair must have been seen three times before the timed region: eager, captured, replayed
    warm_steps = 3 * pipe.depth // math.gcd(n_batches, pipe.depth)

    def timed(on_host, beam, steps, warm):
        for _ in range(max(warm, warm_steps)):
            one_step(on_host, beam)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        # One pair of CUDA events around the K steps.  The host does not wait between steps (a job that captions
        # batch after batch keeps the lanes fed); each step starts with the L2 flush on its first lane's stream,
        # INSIDE the timed region.
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if world > 1:  # device-side rendezvous: every rank's start event sits right behind the same all-reduce
            dist.all_reduce(align)
        t_host0 = time.perf_counter()
        e0.record()
        host_us = []
        for _ in range(steps):
            t0 = time.perf_counter()
            one_step(on_host, beam)
            host_us.append((time.perf_counter() - t0) * 1e6)
        pipe.drain()
        e1.record()
        torch.cuda.synchronize()
        t_host1 = time.perf_counter()
        host_us.sort()
        ms = e0.elapsed_time(e1)
        sys.stderr.write("[bench rank %d] %s beam %d: %.3f ms per step on this rank; host time per step: min %.0f / median %.0f / "
                         "max %.0f us; %s\n" % (rank, "host buffers" if on_host else "resident", beam, ms / steps, host_us[0],
                                                host_us[len(host_us) // 2], host_us[-1], nvml_state(local)))
        if world > 1:
            dist.barrier()
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, (t_host0, t_host1)

    beam = w["beam"]
    other_beam = None if args.no_second_beam else (3 if beam == 1 else 1)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    windows = []
    eng.launch_count(reset=True)
    ms_dev, win = timed(False, beam, args.steps, args.warmup)
    windows.append(win)
    launches = eng.launch_count(reset=True) * args.steps // (args.steps + max(args.warmup, warm_steps))
    ms_host, win = timed(True, beam, args.steps, args.warmup)
    windows.append(win)
    second = None
    if other_beam is not None:
        k2 = max(3, args.steps // 4) if other_beam > beam else args.steps
        ms2_dev, win = timed(False, other_beam, k2, args.warmup)
        windows.append(win)
        ms2_host, win = timed(True, other_beam, k2, args.warmup)
        windows.append(win)
        second = {"beam": other_beam, "value": total_events * k2 / (ms2_dev * 1e-3), "unit": UNIT, "steps": k2,
                  "ms_per_step": ms2_dev / k2,
                  "e2e": {"value": total_events * k2 / (ms2_host * 1e-3), "unit": UNIT, "ms_per_step": ms2_host / k2}}
    clocks = sampler.stop(windows) if rank == 0 else None

    # ---- gathered == single-GPU: one more (untimed) step whose gathered token matrices rank 0 compares, batch by
    # batch, with its own one-GPU decode of the same batch of videos (every rank's batches, rebuilt from seeds)
    gathered_ok = None
    if world > 1:
        keep = []
        one_step(False, beam, keep)
        torch.cuda.synchronize()
        got = {slot: (t.wait()[0] if hasattr(t, "wait") else t).clone() for slot, t in keep}
        torch.cuda.synchronize()
        dist.barrier()
        if rank == 0:
            gathered_ok, checked = True, 0
            solo = model.engine.fork()
            solo.set_concurrency(pipe.depth)  # the lanes' kernel schedule
            for slot in range(n_batches):
                row = 0
                for r in range(world):
                    a, b = plan[r][slot]
                    n_r = ev_counts[r][slot]
                    if b > a:
                        bt = mine[mine_idx.index(slot)] if r == 0 else Batch(build(a, b), dev)
                        tok = solo.caption_device(bt.layout(), bt.d[0], bt.d[1], bt.d[2], beam=beam)[0]
                        same = bool(torch.equal(tok, got[slot][row:row + n_r]))
                        gathered_ok = gathered_ok and same
                        checked += n_r
                        if not same:
                            sys.stderr.write("[bench] gathered tokens of rank %d batch %d differ from the one-GPU decode\n" % (r, slot))
                    row += n_r
                gathered_ok = gathered_ok and row == got[slot].shape[0]
            solo.close()
            gathered_ok = {"equal": gathered_ok, "events_checked": checked,
                           "how": "rank 0 re-decoded every batch of every rank on its own GPU and compared the token ids with the "
                                  "NCCL-gathered matrix, bit for bit"}
        dist.barrier()

    # ---- per-kernel timing of one decode step on a lane (the schedule the timed loop runs), CUDA events on the
    # launching stream inside the library, over the last batch the lane decoded
    lane, bt = pipe.lanes[pipe.depth - 1], mine[0]
    lane.caption_device(bt.lay, bt.d[0], bt.d[1], bt.d[2], beam=beam)
    torch.cuda.synchronize()
    kt = lane.time_step_kernels(beam=beam, iters=20)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    pk, pk_kind = peaks()
    rows = lane.layout.n_events * beam
    V1 = opt.CG_vocab_size + 1
    # Tensor-pipe roofline of the two per-step GEMM launches.  `achieved` counts the FLOPs the launch ISSUES
    # (2*rows*N*K over the k-blocks it walks; the embedding third of the gate product is a table gather and, below
    # 1024 rows, the event/video blocks are hoisted out of the step); `algorithmic` is SURVEY.md 8d's figure
    # (gates 2*2048*[(512+512)+(512+500+512)+(512+512)], logits 2*1536*(V+1) per row-step).
    hoisted = rows < 1024
    k_walked = (8 + 16 + 8) * 64 if hoisted else (16 + 16 + 10) * 64
    issued = {"lstm_gates": 2.0 * rows * 2048 * k_walked, "logit": 2.0 * rows * V1 * 1536}
    algorithmic = {"lstm_gates": 2.0 * rows * 2048 * (1024 + 1524 + 1024), "logit": 2.0 * rows * V1 * 1536}
    cands = ["lstm_gates", "logit"]
    if kt.get("logit_and_next_query", 0.0) > 0.0:
        # greedy decode of >= 1024 rows: the logit launch also carries the next step's attention query GEMM
        # ([rows,512] x [512,512]) as its last tiles; that launch, not the bare logit GEMM, is what the step runs
        issued["logit_and_next_query"] = issued["logit"] + 2.0 * rows * 512 * 512
        algorithmic["logit_and_next_query"] = issued["logit_and_next_query"]
        cands[1] = "logit_and_next_query"
    dom = max(cands, key=lambda k: kt[k])
    sustained = (ms_dev / args.steps) > 5.0  # a step of many back-to-back launches: the sustained peak applies
    peak_tf = pk.get("bf16_tflops_sustained", pk["bf16_tflops"]) if sustained else pk["bf16_tflops"]
    ach = issued[dom] / (kt[dom] * 1e-3) / 1e12
    traffic_all = json.load(open(TRAFFIC_FILE)) if os.path.isfile(TRAFFIC_FILE) else {}
    tkey = "%s/%s/beam%d/rows%d" % (args.workload, args.precision, beam, rows)
    tr = traffic_all.get(tkey, {})
    roofline = {"bound": "tensor", "kernel": "gemm_tc%s_kernel<%s>" % ("" if hoisted else "2", dom),
                "achieved": ach,
                "peak": peak_tf, "unit": "TFLOP/s", "frac": ach / peak_tf, "traffic": tr.get(dom),
                "peak_source": pk_kind + (", sustained" if sustained else ", burst"), "rows": rows,
                "ms_per_launch": kt[dom], "step_kernel_ms": kt,
                "flops_per_launch": {"issued": issued, "algorithmic": algorithmic},
                "per_kernel_tflops_issued": {k: issued[k] / (kt[k] * 1e-3) / 1e12 for k in issued},
                "per_kernel_frac_issued": {k: issued[k] / (kt[k] * 1e-3) / 1e12 / peak_tf for k in issued},
                "traffic_key": tkey, "traffic_source": os.path.relpath(TRAFFIC_FILE, ROOT) if tr else None,
                "note": "FLOPs 2*M*N*K; the %s mode issues %d tcgen05 product(s) per k-step"
                        % (args.precision, {"fp32": 3, "bf16w": 2, "f16": 1}[args.precision])}
    # Attention against HBM.  Algorithmic bytes per launch = every clip row that lies inside a proposal ONCE
    # (clip features + projected table; overlapping proposals of a video re-read them from L2, which is not HBM
    # traffic: SURVEY.md 8d's caveat) + per row the query in and the attended feature out.  The summed-window
    # figure and the measured dram / L2 bytes of the ncu capture stand beside it.
    el = lane.attention_bytes_per_element()
    uniq = bt.unique_window_clips() * (500 + 512) * el + rows * (512 * 4 + 500 * 2)
    summed = sum(e - s for _, s, e in bt.windows) * beam * (500 + 512) * el + rows * (512 * 4 + 500 * 2)
    att_s = kt["attention"] * 1e-3
    roofline["attention"] = {"bound": "hbm", "achieved": uniq / att_s / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s",
                             "frac": uniq / att_s / 1e9 / pk["hbm_gbs"], "algorithmic_bytes": uniq,
                             "summed_window_bytes": summed, "summed_window_gbs": summed / att_s / 1e9,
                             "traffic": tr.get("attention"), "l2_bytes": tr.get("attention_l2"),
                             "bytes_per_element": el, "ms_per_launch": kt["attention"],
                             "kernel": "attention_f16_kernel" if el == 2 else "attention_v2_kernel"}
    # its second bound: one reciprocal per (row, clip, hidden unit) on the special-function pipe, 16 per clock and SM
    sfu_ops = sum(e - s for _, s, e in bt.windows) * beam * 512
    sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
    sfu_rate = 16.0 * sm_count * (clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965) * 1e6
    roofline["attention"]["sfu"] = {"ops": sfu_ops, "floor_us": sfu_ops / sfu_rate * 1e6,
                                    "frac_of_launch": sfu_ops / sfu_rate / att_s}

    line = {"metric": METRIC, "value": total_events * args.steps / (ms_dev * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": {"fp32": "bf16x3 (fp32-grade)", "bf16w": "bf16", "f16": "fp16"}[args.precision],
            "data": "synthetic", "config": config,
            "run": {"precision": args.precision, "in_flight": pipe.depth, "batch_videos": args.batch_videos,
                    "batches_per_rank_per_step": n_batches, "events_rank0": n_events, "rows_per_batch": rows,
                    "ms_per_decode_step_of_a_batch": ms_dev / args.steps / n_batches / S,
                    "precision_note": {"f16": "every tensor-core operand one fp16 plane (11-bit mantissa >= bf16's 8; a "
                                              "bf16-representable checkpoint is exact in fp16), fp32 accumulation, state and softmax",
                                       "bf16w": "activations bf16 hi+lo planes, weights bf16",
                                       "fp32": "bf16 hi+lo planes on both operands (3 products): fp32-grade"}[args.precision]},
            "clocks": clocks,
            "e2e": {"value": total_events * args.steps / (ms_host * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": sum(b.h2d_bytes() for b in mine),
                    "d2h_bytes_per_step": int(n_events * S * 8 + (n_events * 4 if beam > 1 else 0)),
                    "ms_per_step": ms_host / args.steps,
                    "note": "per batch: BatchLayout from the python proposal lists, H2D of c3d / gathered proposal states / lda / "
                            "index arrays, encode, decode, D2H of tokens + log-probs; bytes are rank 0's"},
            "gpu_launches": int(launches),
            "gathered_equals_single_gpu": gathered_ok if world > 1 else "n/a (one rank: nothing is gathered)",
            "roofline": roofline}
    if second is not None:
        line["beam%d" % other_beam] = second
    if not args.no_cpu_baseline and world == 1:
        sample = min(args.cpu_sample or CPU_SAMPLE[args.workload], w["n_videos"])
        reps = 5 if args.workload in ("cfg2", "cfg5") else 3
        sub = [synth.make_video(i, T=w["T"], n_events=w["n_events"], style=w["style"]) for i in range(sample)]
        v, sec, kind = cpu_reference_run(sub, sd_full, beam, reps=reps, warm=1)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": kind,
                                "sample": "first %d videos (%d events) of the workload, one video per call, "
                                          "median of %d passes after 1 warm-up (%.2f s per pass)"
                                          % (sample, sum(x.n_events for x in sub), reps, sec)}
        if other_beam is not None and other_beam > 1 and args.workload in ("cfg2", "cfg5"):
            sub3 = sub[:8]
            v3, sec3, kind3 = cpu_reference_run(sub3, sd_full, other_beam, reps=3, warm=1)
            line["beam%d" % other_beam]["cpu_baseline"] = {
                "value": v3, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": kind3,
                "sample": "first 8 videos (%d events), the oracle's beam loop driving the reference's get_logprobs_state "
                          "(the reference has no beam search), median of 3 passes (%.2f s per pass)"
                          % (sum(x.n_events for x in sub3), sec3)}
    _RESULT.append(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
