"""Per-kernel counts of the SASS instructions that prove the Blackwell paths (B200_PROFILING.md): UTCHMMA (tcgen05.mma),
LDTM (tcgen05.ld), UTMALDG (TMA tensor load), UBLKCP (cp.async.bulk), UTCBAR (tcgen05.commit), SYNCS (mbarrier),
MUFU.  Runs on the authoring machine (cuobjdump only):  python tools/sass_counts.py [tag]  -> profiles/<tag>_sass_counts.md"""
import collections
import re
import subprocess
import sys

tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
so = "echr_b200/libechr_b200.so"
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
names = subprocess.run(["cu++filt"], input="\n".join(re.findall(r"Function : (\S+)", out)), capture_output=True, text=True).stdout.split("\n")
mn = ("UTCHMMA", "LDTM", "UTMALDG", "UBLKCP", "UTCBAR", "SYNCS", "MUFU", "LDGSTS", "LDG", "STG", "ATOM", "RED")
rows = []
cur = None
counts = None
it = iter(names)
for line in out.split("\n"):
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        if cur is not None:
            rows.append((cur, counts))
        cur, counts = next(it), collections.Counter()
        continue
    if cur is None:
        continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        for k in mn:
            if op.startswith(k):
                counts[k] += 1
        counts["total"] += 1
if cur is not None:
    rows.append((cur, counts))


def short(n):
    i = n.rfind("(")
    n = n[:i] if i > 0 else n
    return n.replace("echr::", "").replace("void ", "").replace("(bool)1", "true").replace("(bool)0", "false")


with open("profiles/%s_sass_counts.md" % tag, "w") as f:
    f.write("# SASS instruction counts per kernel of `%s` (`cuobjdump -sass`, sm_100a)\n\n" % so)
    f.write("UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = cp.async.bulk.tensor (TMA), UBLKCP = cp.async.bulk, UTCBAR = tcgen05.commit, "
            "SYNCS = mbarrier ops, LDGSTS = cp.async (the staged LSTM epilogue).  Template arguments: gemm_tc_kernel<A_TERMS, W_TERMS, BLOCK_N, "
            "EPI (0 store, 1 LSTM, 2 logits), TOPK, SPLIT>, gemm_tc2_kernel<A_TERMS, W_TERMS, EPI, TOPK>, pair_mlp_kernel<A_TERMS, W_TERMS>.\n\n")
    f.write("| kernel | instr | " + " | ".join(mn) + " |\n|---|---|" + "---|" * len(mn) + "\n")
    for n, c in sorted(rows, key=lambda r: short(r[0])):
        f.write("| `%s` | %d | %s |\n" % (short(n), c["total"], " | ".join(str(c[k]) for k in mn)))
    tot = collections.Counter()
    for _, c in rows:
        tot.update(c)
    f.write("\nTotals over %d kernels: %s\n" % (len(rows), ", ".join("%s %d" % (k, tot[k]) for k in mn)))
print("wrote profiles/%s_sass_counts.md (%d kernels)" % (tag, len(rows)))
