"""Join an ncu source-page CSV (per-SASS-instruction stall samples) with nvdisasm line info, so samples
can be read per CUDA source line and per kernel phase (the ncu CSV export carries no line numbers).
usage: ncu_by_line.py <report.ncu-rep> <cubin> <mangled-name substring> <ncu kernel-name substring> [top]"""
import collections
import csv
import os
import re
import subprocess
import sys

rep, cubin, ksub, nsub = sys.argv[1:5]
top = int(sys.argv[5]) if len(sys.argv) > 5 else 30
sass = subprocess.run(["nvdisasm", "-g", "-c", cubin], stdout=subprocess.PIPE, text=True).stdout.splitlines()
cur_fn, cur_line, table = None, None, {}
for ln in sass:
    m = re.match(r"^\.text\.(\S+):", ln)
    if m:
        cur_fn = m.group(1)
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur_line = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"^\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m and cur_fn and ksub in cur_fn:
        table[int(m.group(1), 16)] = (cur_line, m.group(2).strip())
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(out.splitlines()))
KEYS = ("# Samples", "Instructions Executed", "stall_barrier", "stall_long_sb", "stall_short_sb", "stall_mio",
        "stall_wait", "stall_math", "stall_not_selected", "stall_no_inst", "stall_lg", "stall_dispatch",
        "stall_selected", "stall_branch_resolving", "L1 Wavefronts Shared Excessive", "L1 Wavefronts Shared")
hdr, want, base, done = None, False, None, False
agg = collections.defaultdict(collections.Counter)
opagg = collections.defaultdict(collections.Counter)
tot = collections.Counter()
for r in rows:
    if r and r[0] == "Kernel Name":
        if want and base is not None:
            done = True
        want = (nsub in r[1]) and not done
        hdr, base = None, None
        continue
    if r and r[0] == "Address":
        hdr = r
        continue
    if not hdr or len(r) != len(hdr) or not want:
        continue
    d = dict(zip(hdr, r))
    addr = int(d["Address"], 16)
    if base is None:
        base = addr
    line, ins = table.get(addr - base, ((None, -1), d["Source"].strip()))
    op = d["Source"].strip().split()[0] if d["Source"].strip() else "?"
    if op.startswith("@"):
        op = d["Source"].strip().split()[1]
    op = op.split(".")[0]
    for k in KEYS:
        try:
            v = float(d[k])
        except (KeyError, ValueError):
            v = 0.0
        agg[line][k] += v
        opagg[op][k] += v
        tot[k] += v
print("total samples %d, warp instructions %d" % (tot["# Samples"], tot["Instructions Executed"]))
print("stall totals (%% of samples):", {k[6:]: round(100 * v / tot["# Samples"], 1) for k, v in tot.items() if k.startswith("stall")})
print("smem wavefronts %d, excessive %d" % (tot["L1 Wavefronts Shared"], tot["L1 Wavefronts Shared Excessive"]))
print("\nopcode mix (share of executed warp instructions):")
for op, c in sorted(opagg.items(), key=lambda kv: -kv[1]["Instructions Executed"])[:14]:
    print("  %-10s %6.2f%% inst  %6.2f%% samples" % (op, 100 * c["Instructions Executed"] / tot["Instructions Executed"], 100 * c["# Samples"] / tot["# Samples"]))
print("\ntop source lines:")
for line, c in sorted(agg.items(), key=lambda kv: -kv[1]["# Samples"])[:top]:
    print("%5.2f%% smp %5.2f%% inst  %-24s bar %6d lsb %6d ssb %6d mio %6d wait %6d disp %6d nsel %6d lg %5d | xs-wf %d" % (
        100 * c["# Samples"] / tot["# Samples"], 100 * c["Instructions Executed"] / tot["Instructions Executed"],
        "%s:%s" % line, c["stall_barrier"], c["stall_long_sb"], c["stall_short_sb"], c["stall_mio"], c["stall_wait"],
        c["stall_dispatch"], c["stall_not_selected"], c["stall_lg"], c["L1 Wavefronts Shared Excessive"]))
# ---- phase view: fused_kernel.cuh line ranges delimited by its section comments
src = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tensordiffeq_b200", "csrc", "fused_kernel.cuh")).read().splitlines()
marks = [(i, m.group(1).strip("= -")[:64]) for i, l in enumerate(src, 1) for m in [re.search(r"// (=+ .* =+|---- .*)$", l)] if m]
ph, phi = collections.Counter(), collections.Counter()
for (f, line), c in agg.items():
    name = "(inlined: %s)" % f
    if f == "fused_kernel.cuh":
        name = "prologue"
        for ml, nm in marks:
            if line >= ml:
                name = "%4d %s" % (ml, nm)
    ph[name] += c["# Samples"]
    phi[name] += c["Instructions Executed"]
print("\nphase view (samples ~ warp time):")
for k, v in sorted(ph.items()):
    print("  %5.2f%% time  %5.2f%% inst  %s" % (100 * v / tot["# Samples"], 100 * phi[k] / tot["Instructions Executed"], k))
