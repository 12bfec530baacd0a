#!/usr/bin/env python
"""Benchmark of the ECHR caption-generation hot path (BASELINE.json metric: events captioned / s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2] [--beam 1] [--precision f16] [--in-flight 4]
    python bench.py --impl reference ...      # the reference algorithm's CPU path (oracle port) on the host cores

One "step" = one pass of the hot path (context assembly + relation encoder + 30-token decode) over
one batch of synthetic videos.  Videos are independent, so with N GPUs every rank captions its own
`workload` batch (weak scaling) and the only collective is one all_gather of the token ids per step.
Prints ONE JSON line on rank 0.
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

import torch  # noqa: E402

from echr_b200 import synth  # noqa: E402
from echr_b200.config import echr_opt  # noqa: E402

METRIC = "events_captioned_per_sec"
UNIT = "events/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(synth.WORKLOADS))
    ap.add_argument("--videos", type=int, default=None, help="override the workload's video count (per GPU)")
    ap.add_argument("--beam", type=int, default=None)
    ap.add_argument("--precision", default="f16", choices=["fp32", "bf16w", "f16"])
    ap.add_argument("--cpu-sample", type=int, default=None, help="videos in the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--in-flight", type=int, default=4,
                    help="batches kept in flight per GPU (CaptionPipeline lanes); 1 = one blocking batch at a time")
    return ap.parse_args()


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


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
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def nvml_state(index):
    """One NVML reading of this rank's GPU (diagnosis line on stderr under torchrun)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        return "sm %d MHz, mem %d MHz, %.0f W, event reasons 0x%x" % (
            pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_MEM),
            pynvml.nvmlDeviceGetPowerUsage(h) / 1e3, pynvml.nvmlDeviceGetCurrentClocksEventReasons(h))
    except Exception as e:  # noqa: BLE001
        return "nvml unavailable (%s)" % type(e).__name__


def cpu_reference_run(vids, sd, beam, reps, warm):
    """The reference algorithm's own work shape on the host cores (oracle port, torch CPU fp32):
    one video per call, looped.  Returns (events/s, seconds per pass)."""
    from oracle import echr_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    torch.set_grad_enabled(False)
    n_events = sum(v.n_events for v in vids)

    def one_pass():
        for v in vids:
            if beam > 1:
                O.caption_video_beam(sd, v.tap_feats, v.c3d_feats, v.lda_feats, v.ind_select_list,
                                     v.soi_select_list, beam=beam)
            else:
                O.caption_video(sd, v.tap_feats, v.c3d_feats, v.lda_feats, v.ind_select_list, v.soi_select_list)

    for _ in range(warm):
        one_pass()
    times = []
    for _ in range(reps):
        t0 = time.perf_counter()
        one_pass()
        times.append(time.perf_counter() - t0)
    times.sort()
    med = times[len(times) // 2]
    return n_events / med, med


def workload_config(args, w, n_events, n_gpus):
    return {"workload": "%s: %d videos x T=%d clips, %d events (%s proposals), beam=%d, vocab 5747, max_len 30"
                        % (args.workload, w["n_videos"], w["T"], n_events, w["style"], w["beam"]),
            "videos_per_gpu": w["n_videos"], "events_per_gpu": n_events, "beam": w["beam"],
            "parallelism": "video-sharded dp%d, one all_gather of token ids per step" % n_gpus}


def main_reference(args, rank, world):
    if rank != 0:
        return 0
    opt = echr_opt()
    sd = synth.init_state_dict(opt, seed=0)
    vids, w = synth.make_workload(args.workload, n_videos=args.videos)
    if args.beam is not None:
        w["beam"] = args.beam
    sample = args.cpu_sample or {"cfg1": 1, "cfg2": 64, "cfg3": 1, "cfg4": 1, "cfg5": 64}[args.workload]
    sub = vids[:sample]
    n_ev = sum(v.n_events for v in sub)
    value, sec = cpu_reference_run(sub, sd, w["beam"], reps=max(1, args.steps), warm=max(0, args.warmup))
    cores = os.cpu_count() or 1
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, w, sum(v.n_events for v in vids), world),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "first %d videos (%d events) of the workload, one video per call, median of %d passes"
                                       % (sample, n_ev, max(1, args.steps))},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    _RESULT.append(json.dumps(line))
    return 0


def main():
    args = parse()
    # stdout carries exactly one JSON line: anything a library prints there meanwhile (NCCL's version banner, ...)
    # is sent to stderr; the real stdout comes back for the final print
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        return _main(args)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
        if _RESULT:
            print(_RESULT[0], flush=True)


_RESULT = []


def _main(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return main_reference(args, rank, world)

    import torch.distributed as dist
    from echr_b200.caption_generator import CaptionGenerator
    from echr_b200.engine import BatchLayout, CaptionPipeline
    from echr_b200.parallel import exchange_counts, gather_tokens

    torch.set_grad_enabled(False)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # stdout carries exactly one JSON line: keep NCCL's banner (NCCL_DEBUG=VERSION/INFO prints to stdout) out of it
        os.environ["NCCL_DEBUG"] = os.environ.get("ECHR_NCCL_DEBUG", "WARN")
        dist.init_process_group("nccl", device_id=dev)

    opt = echr_opt()
    sd_full = synth.init_state_dict(opt, seed=0)
    sd = synth.round_state_dict_bf16(sd_full) if args.precision != "fp32" else sd_full
    model = CaptionGenerator(opt, precision=args.precision)
    model.load_state_dict(sd, strict=True)
    model = model.to(dev)
    eng = model.engine

    # this rank's shard: its own block of videos (seeds disjoint across ranks)
    w0 = synth.WORKLOADS[args.workload]
    nv = args.videos or w0["n_videos"]
    vids, w = synth.make_workload(args.workload, n_videos=nv, first_seed=rank * nv,
                                  bf16_features=(args.precision != "fp32"))
    if args.beam is not None:
        w["beam"] = args.beam
    beam = w["beam"]
    lay = BatchLayout([v.c3d_feats.shape[0] for v in vids], [v.ind_select_list for v in vids],
                      [v.soi_select_list for v in vids])
    n_events = lay.n_events
    h_c3d = torch.cat([v.c3d_feats for v in vids]).contiguous().pin_memory()
    h_tap = torch.cat([v.tap_feats for v in vids]).contiguous().pin_memory()
    h_lda = torch.stack([v.lda_feats for v in vids]).contiguous().pin_memory()
    d_c3d, d_tap, d_lda = h_c3d.to(dev), h_tap.to(dev), h_lda.to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    counts = exchange_counts(n_events, dev) if world > 1 else None
    align = torch.zeros(1, device=dev)
    no_gather = os.environ.get("ECHR_BENCH_NO_GATHER") == "1"  # diagnosis only: the line then says so

    # the public call of a captioning job: batches submitted one after the other to a CaptionPipeline, which
    # keeps `--in-flight` of them on the GPU (one workspace + stream per lane, one packed model)
    pipe = CaptionPipeline(eng, depth=max(1, args.in_flight))

    def gather(res):
        return (gather_tokens(res[0], counts=counts),) + tuple(res[1:])

    def gather_host(res):
        return (gather_tokens(res[0].to(dev, non_blocking=True), counts=counts),) + tuple(res[1:])

    def step_device():
        return pipe.submit(lay, d_c3d, d_tap, d_lda, beam=beam, before=flush.zero_, after=gather if world > 1 and not no_gather else None)

    def step_host():
        return pipe.submit(lay, h_c3d, h_tap, h_lda, beam=beam, before=flush.zero_, after=gather_host if world > 1 and not no_gather else None)

    def timed(fn, steps, warm):
        for _ in range(max(warm, 2 * pipe.depth)):  # every lane sees the shape twice (eager, then captured)
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        # One pair of CUDA events around the K steps.  The host does not wait between steps (a job that captions
        # batch after batch keeps the lanes fed); each step starts with the L2 flush on its lane's stream, so
        # the flush is INSIDE the timed region (it overlaps the other lane's decode when in_flight > 1).
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if world > 1:  # device-side rendezvous: every rank's start event sits right behind the same all-reduce,
            dist.all_reduce(align)  # so a rank whose host left the barrier late does not stretch the others' region
        e0.record()
        tickets, host_us = [], []
        for _ in range(steps):
            t0 = time.perf_counter()
            tickets.append(fn())
            host_us.append((time.perf_counter() - t0) * 1e6)
        host_us.sort()
        sys.stderr.write("[bench rank %d] host time per submit: min %.0f / median %.0f / max %.0f us\n"
                         % (rank, host_us[0], host_us[len(host_us) // 2], host_us[-1]))
        pipe.drain()
        e1.record()
        torch.cuda.synchronize()
        for t in tickets[-pipe.depth:]:
            t.wait()
        ms = e0.elapsed_time(e1)
        if world > 1:
            sys.stderr.write("[bench rank %d] %.4f ms per step on this rank; %s\n" % (rank, ms / steps, nvml_state(local)))
            dist.barrier()
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    sampler = ClockSampler(local)
    if rank == 0 and os.environ.get("ECHR_BENCH_NO_SAMPLER") != "1":  # (diagnosis switch; the line then has no clocks)
        sampler.start()
    eng.launch_count(reset=True)
    ms_dev = timed(step_device, args.steps, args.warmup)
    launches = eng.launch_count(reset=True) // max(1, args.steps + max(args.warmup, 2 * pipe.depth)) * args.steps
    ms_host = timed(step_host, args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None

    # per-kernel timing of one decode step (CUDA events on the launching stream, inside the library)
    kt = eng.time_step_kernels(beam=beam, iters=20)

    total_events = n_events * world
    if world > 1:
        t = torch.tensor([n_events], dtype=torch.int64, device=dev)
        dist.all_reduce(t)
        total_events = int(t.item())
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    pk, pk_kind = peaks()
    rows = n_events * beam
    V1, S = opt.CG_vocab_size + 1, opt.CG_seq_length
    # dominant kernel: the slower of the two GEMM launches of a step.  Algorithmic FLOPs per row-step
    # (constant-context terms hoisted): gates 2*2048*[(512+512)+(512+500+512)+(512+512)],
    # logits 2*1536*(V+1).
    flops = {"lstm_gates": 2.0 * rows * 2048 * (1024 + 1524 + 1024), "logit": 2.0 * rows * V1 * 1536}
    dom = max(("lstm_gates", "logit"), key=lambda k: kt[k])
    ach = flops[dom] / (kt[dom] * 1e-3) / 1e12
    peak_tf = pk.get("bf16_tflops_sustained", pk["bf16_tflops"])
    # DRAM bytes per launch of that kernel from the committed `ncu --set full` capture of this command
    # (profiles/r01_traffic.json; cold-cache replays, so the weights count once per launch)
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json")))
        key = "%s/%s/beam%d" % (args.workload if args.videos is None else "%s-%d" % (args.workload, args.videos),
                                args.precision, beam)
        traffic = tj.get(key, {}).get(dom)
    except Exception:
        traffic = None
    roofline = {"bound": "tensor", "kernel": "gemm_tc_kernel<%s>" % dom, "achieved": ach, "peak": peak_tf,
                "unit": "TFLOP/s", "frac": ach / peak_tf, "traffic": traffic, "peak_source": pk_kind + " (sustained)",
                "ms_per_launch": kt[dom], "step_kernel_ms": kt,
                "per_kernel_tflops": {k: flops[k] / (kt[k] * 1e-3) / 1e12 for k in flops},
                "note": "algorithmic FLOPs 2*M*N*K of one product; the %s mode issues %d tcgen05 products per k-step"
                        % (args.precision, {"fp32": 3, "bf16w": 2, "f16": 1}[args.precision])}
    # attention kernel against HBM (2024*L + 4048 bytes per event-step at bf16, fp32 clip doubles the window)
    el = 4  # clip features and the e^{2P} table are read as fp32 in every precision
    win = sum((s[1] - s[0]) for v in vids for s in v.soi_select_list) * beam
    att_bytes = win * (500 + 512) * el + rows * (512 * 4 + 500 * 4)
    roofline["attention"] = {"bound": "hbm", "achieved": att_bytes / (kt["attention"] * 1e-3) / 1e9,
                             "peak": pk["hbm_gbs"], "unit": "GB/s",
                             "frac": att_bytes / (kt["attention"] * 1e-3) / 1e9 / pk["hbm_gbs"]}

    line = {"metric": METRIC, "value": total_events * args.steps / (ms_dev * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": {"fp32": "bf16x3 (fp32-grade)", "bf16w": "bf16", "f16": "fp16"}[args.precision],
            "data": "synthetic",
            "config": dict(workload_config(args, w, n_events, world), precision=args.precision,
                           l2_flush_between_steps=True, in_flight=pipe.depth,
                           precision_note={"f16": "every tensor-core operand one fp16 plane (11-bit mantissa >= bf16's 8; a "
                                                  "bf16-representable checkpoint is exact in fp16), fp32 accumulation, state and "
                                                  "softmax; tokens equal the fp32-grade path except at top-2 gaps < 1e-4 "
                                                  "(profiles/r01_precision_token_agreement.txt)",
                                           "bf16w": "activations bf16 hi+lo planes, weights bf16: bit-identical to the fp32-grade "
                                                    "path on a bf16-representable checkpoint",
                                           "fp32": "bf16 hi+lo planes on both operands (3 products): fp32-grade"}[args.precision],
                           **({"DIAGNOSIS_no_token_gather": True} if no_gather and world > 1 else {}),
                           l2_flush="256 MB write at the head of every step, on the step's stream, inside the timed region",
                           ms_per_decode_step=ms_dev / args.steps / S),
            "clocks": clocks,
            "e2e": {"value": total_events * args.steps / (ms_host * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": int(h_c3d.numel() + n_events * 512 + h_lda.numel()) * 4 + int(9 * n_events + 2 * len(vids)) * 4,
                    "d2h_bytes_per_step": int(n_events * S * 8 + (n_events * 4 if beam > 1 else 0)),
                    "ms_per_step": ms_host / args.steps},
            "gpu_launches": int(launches),
            "roofline": roofline}
    if not args.no_cpu_baseline and world >= 1:
        # bounded sample sized for ~10-20 s of CPU work: the whole cfg2 workload (64 videos) takes ~2 s per pass
        sample = args.cpu_sample or {"cfg1": 1, "cfg2": 64, "cfg3": 1, "cfg4": 1, "cfg5": 64}[args.workload]
        reps = 5 if args.workload in ("cfg2", "cfg5") else 3
        sub = [synth.make_video(i, T=w["T"], n_events=w0["n_events"], style=w["style"]) for i in range(sample)]
        v, sec = cpu_reference_run(sub, sd_full, beam, reps=reps, warm=1)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "port",
                                "sample": "first %d videos (%d events) of the workload, one video per call, "
                                          "median of %d passes after 1 warm-up (%.2f s per pass)"
                                          % (sample, sum(x.n_events for x in sub), reps, sec)}
    _RESULT.append(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
