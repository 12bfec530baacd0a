"""Turns ncu output brought back in gpurun_out/ into the tracked summaries under profiles/.

    python tools/summarize_profile.py launches <launches.csv> <out.md> "<command the list was taken from>"
    python tools/summarize_profile.py table <raw page .csv> <out_prefix> "<title>"
    python tools/summarize_profile.py full <report.ncu-rep> <out_prefix> "<title>"     (needs ncu on this machine)
"""
import collections
import csv
import re
import subprocess
import sys

KEYS = ["Kernel Name", "launch__grid_size", "launch__block_size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "lts__t_bytes.sum", "l1tex__m_xbar2l1tex_read_bytes.sum", "lts__t_sector_hit_rate.pct",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic"]


def short(name):
    name = re.sub(r"\(.*", "", name).replace("echr::", "").replace("void ", "")
    return name[:90]


def launches(path, out, cmd):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    iK, iV, iG = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        try:
            ns = float(r[iV].replace(",", ""))
        except ValueError:
            continue
        a = agg.setdefault((short(r[iK]), r[iG]), [0, 0.0])
        a[0] += 1
        a[1] += ns
    tot = sum(a[1] for a in agg.values())
    with open(out, "w") as f:
        f.write("# ncu launch list of `%s`\n\n" % cmd)
        f.write("`ncu --metrics gpu__time_duration.sum --clock-control none` (per-launch times are cold-cache and serialised: "
                "compare SHARES, not absolutes).  %d launches, %.2f ms in total.\n\n" % (sum(a[0] for a in agg.values()), tot / 1e6))
        f.write("| kernel | grid | launches | total us | avg us | share |\n|---|---|---|---|---|---|\n")
        for (name, grid), (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("| `%s` | %s | %d | %.1f | %.2f | %.1f%% |\n" % (name, grid, n, ns / 1e3, ns / n / 1e3, 100 * ns / tot))
        f.write("\nTemplate arguments: gemm_tc_kernel<A_TERMS, W_TERMS, BLOCK_N, EPI (0 store, 1 LSTM, 2 logits), TOPK, SPLIT>; "
                "gemm_tc2_kernel<A_TERMS, W_TERMS, EPI, TOPK, ATT>; pair_mlp_kernel<A_TERMS, W_TERMS>.\n")
    print("wrote", out)


def full(rep, prefix, title):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    open(prefix + "_raw.csv", "w").write(raw)
    rr = list(csv.reader(raw.splitlines()))
    h, u = rr[0], rr[1]
    with open(prefix + ".md", "w") as f:
        f.write("# %s\n\n`ncu --set full --clock-control none` (cold-cache replays: ncu flushes L2 between passes, so DRAM bytes include "
                "operands that are L2-resident inside the real decode graph).  Raw page: %s_raw.csv\n\n" % (title, prefix.split("/")[-1]))
        keys = [k for k in KEYS if k in h]
        f.write("| " + " | ".join(k.split(".")[0] for k in keys) + " |\n|" + "---|" * len(keys) + "\n")
        for r in rr[2:]:
            f.write("| " + " | ".join((short(r[h.index(k)]) if k == "Kernel Name" else r[h.index(k)][:24] + " " + u[h.index(k)]) for k in keys) + " |\n")
    print("wrote", prefix + ".md")


def table(raw_csv, prefix, title):
    """The raw page of an `ncu --set full` capture exported on the GPU box (tools/gpu_round.sh ncufull) -> one row per
    (kernel, grid): the median launch of each group, the columns the roofline arguments use, the three largest warp-stall
    reasons (warps stalled per issue-active cycle).  Also writes the selected columns of EVERY launch to <prefix>_raw.csv."""
    rr = list(csv.reader(open(raw_csv)))
    h, u = rr[0], rr[1]
    ci = {k: i for i, k in enumerate(h)}
    cols = ["Kernel Name", "launch__grid_size", "launch__block_size", "gpu__time_duration.sum", "dram__bytes_read.sum",
            "dram__bytes_write.sum", "l1tex__m_xbar2l1tex_read_bytes.sum", "lts__t_sector_hit_rate.pct",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "dram__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
            "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum"]
    cols = [c for c in cols if c in ci]
    stall = [k for k in h if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("per_issue_active.ratio")]
    with open(prefix + "_raw.csv", "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(cols + stall)
        w.writerow([u[ci[c]] for c in cols + stall])
        for r in rr[2:]:
            w.writerow([short(r[ci[c]]) if c == "Kernel Name" else r[ci[c]] for c in cols + stall])

    def num(r, k):
        try:
            return float(r[ci[k]].replace(",", ""))
        except (ValueError, KeyError):
            return float("nan")
    groups = collections.OrderedDict()
    for r in rr[2:]:
        groups.setdefault((short(r[ci["Kernel Name"]]), r[ci["launch__grid_size"]]), []).append(r)
    with open(prefix + ".md", "w") as f:
        f.write("# %s\n\n`ncu --set full --clock-control none` of `python tools/profile_targets.py` (3-token captions; cold-cache, serialised "
                "replays: DRAM bytes include operands that stay in L2 inside the real decode graph, durations are longer than in "
                "the graph).  One row per (kernel, grid): the median launch.  `l2->sm` = l1tex__m_xbar2l1tex_read_bytes; `xu` = "
                "special-function pipe; stalls = warps stalled per issue-active cycle, three largest reasons.  "
                "Every launch: %s_raw.csv\n\n" % (title, prefix.split("/")[-1]))
        f.write("| kernel | grid | launches | us | dram rd MB | dram wr MB | l2->sm MB | tensor % | xu % | warps % | sm % | regs | stalls |\n"
                "|---|---|---|---|---|---|---|---|---|---|---|---|---|\n")
        for (name, g), rows in groups.items():
            r = rows[len(rows) // 2]
            st = sorted(((num(r, k), k.split("stalled_")[1].split("_per_issue")[0]) for k in stall), reverse=True)[:3]
            f.write("| `%s` | %s | %d | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %d | %s |\n" % (
                name, g, len(rows), num(r, "gpu__time_duration.sum"), num(r, "dram__bytes_read.sum"), num(r, "dram__bytes_write.sum"),
                num(r, "l1tex__m_xbar2l1tex_read_bytes.sum"), num(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
                num(r, "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"), num(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
                num(r, "sm__throughput.avg.pct_of_peak_sustained_elapsed"), int(num(r, "launch__registers_per_thread")),
                ", ".join("%s %.1f" % (b, a) for a, b in st)))
        f.write("\nUnits as exported by ncu: %s for durations, %s for bytes.\n" % (u[ci["gpu__time_duration.sum"]], u[ci["dram__bytes_read.sum"]]))
    print("wrote", prefix + ".md")


if __name__ == "__main__":
    mode = sys.argv[1]
    if mode == "launches":
        launches(sys.argv[2], sys.argv[3], sys.argv[4])
    elif mode == "table":
        table(sys.argv[2], sys.argv[3], sys.argv[4])
    else:
        full(sys.argv[2], sys.argv[3], sys.argv[4])
