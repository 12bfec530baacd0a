"""Turns gpurun_out/prof_launches.csv (tools/gpu_profile.sh) and gpurun_out/prof_step.ncu-rep (tools/gpu_profile_full.sh) into the
tracked summaries under profiles/:  python tools/summarize_profile.py <tag>"""
import collections
import csv
import subprocess
import sys

tag = sys.argv[1] if len(sys.argv) > 1 else "r01_final"
rows = [r for r in csv.reader(open("gpurun_out/prof_launches.csv")) if len(r) > 5]
hdr = rows[0]
iK, iV, iG = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size")
agg = collections.OrderedDict()
seq = []
for r in rows[1:]:
    try:
        ns = float(r[iV].replace(",", ""))
    except ValueError:
        continue
    name = r[iK]
    key = (name[:64], r[iG])
    a = agg.setdefault(key, [0, 0.0])
    a[0] += 1
    a[1] += ns
    seq.append((name, r[iG], ns))
tot = sum(a[1] for a in agg.values())
with open("profiles/%s_launch_shares_cfg2_f16.md" % tag, "w") as f:
    f.write("# ncu launch list of `python bench.py --steps 2 --warmup 2 --no-cpu-baseline` (cfg2, f16, 3 batches in flight)\n\n")
    f.write("`ncu --metrics gpu__time_duration.sum --clock-control none -c 2000` (cold-cache, serialised: compare shares). "
            "Raw list: %s_launches_cfg2_f16.csv\n\n" % tag)
    f.write("| kernel | grid | launches | total us | avg us | share |\n|---|---|---|---|---|---|\n")
    for (name, grid), (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write("| `%s` | %s | %d | %.1f | %.2f | %.1f%% |\n" % (name, grid, n, ns / 1e3, ns / n / 1e3, 100 * ns / tot))
    f.write("\nTemplate args of gemm_tc_kernel: <A_TERMS, W_TERMS, BLOCK_N, EPI (0 store, 1 LSTM, 2 logits), TOPK, SPLIT (cluster size)>.\n")
subprocess.check_call("cp gpurun_out/prof_launches.csv profiles/%s_launches_cfg2_f16.csv" % tag, shell=True)

raw = subprocess.run(["ncu", "-i", "gpurun_out/prof_step.ncu-rep", "--page", "raw", "--csv"], capture_output=True, text=True).stdout
open("profiles/%s_ncu_step_kernels_raw.csv" % tag, "w").write(raw)
rr = list(csv.reader(raw.splitlines()))
h, u = rr[0], rr[1]
keys = ["Kernel Name", "launch__grid_size", "launch__cluster_dim_x", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "lts__t_sector_hit_rate.pct",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "sm__cycles_active.avg"]
with open("profiles/%s_ncu_step_kernels.md" % tag, "w") as f:
    f.write("# `ncu --set full --clock-control none` of the decode-step kernels (cfg2, f16, 251 rows)\n\n"
            "Cold-cache replays (ncu flushes L2 between passes): DRAM bytes include the weights, which stay in L2 "
            "inside the real decode graph.  Raw page: %s_ncu_step_kernels_raw.csv\n\n" % tag)
    f.write("| " + " | ".join(k.split(".")[0] for k in keys) + " |\n|" + "---|" * len(keys) + "\n")
    for r in rr[2:]:
        f.write("| " + " | ".join((r[h.index(k)][:60] + " " + u[h.index(k)]) if k in h else "-" for k in keys) + " |\n")
print("wrote profiles/%s_*" % tag)
