#!/usr/bin/env python3
"""Per-function (kernel + its non-inlined device functions) share of samples / executed instructions and the
top opcodes, from an `ncu --set full --import-source on` report.
  python tools/ncu_by_function.py REPORT.ncu-rep KERNEL_REGEX"""
import collections, csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); h = rows[0]
cols = ["ID", "Kernel Name", "gpu__time_duration.sum", "smsp__inst_executed.sum", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed_pipe_alu.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "dram__bytes_read.sum", "dram__bytes_write.sum"]
for r in rows[2:]:
    print({c.split("__")[-1][:28]: r[h.index(c)][:36] for c in cols if c in h})
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name", "regex:" + kern],
                     capture_output=True, text=True).stdout
fn = hdr = None
tot, tote, ops = collections.Counter(), collections.Counter(), collections.defaultdict(collections.Counter)
for rec in csv.reader(io.StringIO(src)):
    if not rec:
        continue
    if "Source" in rec and "Address" in rec:
        hdr = rec
        ks = [k for k in hdr if k.strip() == "# Samples" or k.startswith("Warp Stall Sampling (All")][0]
        ke = [k for k in hdr if k.startswith("# Warp Instructions Executed") or k == "Instructions Executed"][0]
        continue
    if hdr is None or len(rec) != len(hdr):
        fn = " ".join(rec)[:100]
        continue
    d = dict(zip(hdr, rec))
    try:
        s, e = float(d[ks].replace(",", "")), float(d[ke].replace(",", ""))
    except Exception:
        continue
    tot[fn] += s; tote[fn] += e
    op = d["Source"].split()
    op = op[1] if op and op[0].startswith("@") and len(op) > 1 else (op[0] if op else "?")
    ops[fn][op] += e
S, E = sum(tot.values()), sum(tote.values())
print("kernel regex %s: %d samples, %.4g warp instructions" % (kern, S, E))
for f, s in tot.most_common(14):
    print("%5.1f%% samples %5.1f%% inst  %s" % (100 * s / max(S, 1), 100 * tote[f] / max(E, 1), f))
    print("       ", [(o, round(100 * e / max(tote[f], 1), 1)) for o, e in ops[f].most_common(7)])
