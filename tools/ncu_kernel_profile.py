#!/usr/bin/env python3
"""Per-kernel summary of an `ncu --set full --import-source on` report: headline raw metrics, stall reasons per issue,
opcode mix, and where in the code (4 KB chunks of SASS) the instructions and the stall samples fall.
    python tools/ncu_kernel_profile.py gpurun_out/prof.ncu-rep [--json out.json]"""
import collections
import re
import csv
import io
import json
import subprocess
import sys

rep = sys.argv[1]
out_json = sys.argv[sys.argv.index("--json") + 1] if "--json" in sys.argv else None


def sh(args):
    return subprocess.run(args, check=True, capture_output=True, text=True).stdout


raw = list(csv.reader(io.StringIO(sh(["ncu", "-i", rep, "--page", "raw", "--csv"]))))
hdr, kernels = raw[0], raw[2:]
KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "launch__stack_size", "smsp__inst_executed_op_shared_ld.sum"]
summary = []
for row in kernels:
    d = dict(zip(hdr, row))
    k = {"kernel": d.get("Kernel Name")}
    for key in KEYS:
        if d.get(key) not in (None, "", "n/a"):
            k[key] = d[key]
    stalls = {h.split("issue_stalled_")[1].split("_per_issue")[0]: float(v) for h, v in d.items()
              if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio") and v not in ("", "n/a")}
    k["stall_cycles_per_issue"] = {a: round(b, 3) for a, b in sorted(stalls.items(), key=lambda x: -x[1]) if b >= 0.02}
    summary.append(k)

src = list(csv.reader(io.StringIO(sh(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"]))))
sections, cur, header = [], None, None
for r in src:
    if not r:
        continue
    if r[0] == "Kernel Name":
        cur = []
        sections.append(cur)
    elif r[0] == "Address":
        header = r
    elif cur is not None and header and len(r) == len(header):
        cur.append(dict(zip(header, r)))
# (ncu prints every kernel's section twice when SASS and source views are both available: keep one per kernel)
if len(sections) == 2 * len(summary):
    sections = sections[::2]
lines_of = {}
if "--so" in sys.argv:  # optional: map SASS offsets to CUDA source lines through nvdisasm (needs -lineinfo)
    sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.abspath(__file__)))
    import ncu_hot_lines
    lines_of = ncu_hot_lines.line_map(sys.argv[sys.argv.index("--so") + 1], "kernel")
for k, rows in zip(summary, sections):
    tot = ts = 0.0
    byop, samp, chunks, chs, chv = (collections.Counter() for _ in range(5))
    base = int(rows[0]["Address"], 16)
    for r in rows:
        e, sm = float(r["Instructions Executed"] or 0), float(r["# Samples"] or 0)
        parts = r["Source"].split()
        op = (parts[1] if parts[0].startswith("@") else parts[0]).split(".")[0]
        byop[op] += e
        samp[op] += sm
        tot += e
        ts += sm
        c = (int(r["Address"], 16) - base) // 0x1000
        chunks[c] += e
        chs[c] += sm
        if op == "VABSDIFF4":
            chv[c] += e
    if lines_of:
        name = re.sub(r"[^A-Za-z0-9_]", "", k["kernel"].replace("void ", "").split("(")[0].split("<")[0])
        tmpl = re.search(r"<(\(int\))?(\d+)>|<\(bool\)(\d)>", k["kernel"])
        cands = [fn for fn in lines_of if name in fn]
        if tmpl:
            num = tmpl.group(2) or tmpl.group(3)
            cands = [fn for fn in cands if ("ILi%sE" % num) in fn or ("ILb%sE" % num) in fn] or cands
        if cands:
            mp = lines_of[cands[0]]
            by_line, by_line_e = collections.Counter(), collections.Counter()
            for r in rows:
                hit = mp.get(int(r["Address"], 16) - base)
                by_line[hit[0] if hit else None] += float(r["# Samples"] or 0)
                by_line_e[hit[0] if hit else None] += float(r["Instructions Executed"] or 0)
            k["hot_lines__pct_samples__pct_instr"] = {str(l): [round(100 * v / max(ts, 1), 1), round(100 * by_line_e[l] / tot, 1)]
                                                      for l, v in by_line.most_common(22)}
    k["warp_instructions"] = tot
    k["opcodes_pct_of_instructions__pct_of_stall_samples"] = {op: [round(100 * e / tot, 1), round(100 * samp[op] / max(ts, 1), 1)]
                                                               for op, e in byop.most_common(16)}
    k["code_chunks_4KB__pct_instr__pct_samples__vabs_share"] = {"0x%05x" % (c * 0x1000): [round(100 * chunks[c] / tot, 1), round(100 * chs[c] / max(ts, 1), 1),
                                                                                           round(100 * chv[c] / max(chunks[c], 1))]
                                                                 for c in sorted(chunks) if chunks[c] / tot > 0.01 or chs[c] / max(ts, 1) > 0.01}
for k in summary:
    print("==", k["kernel"])
    for a, b in k.items():
        if a != "kernel":
            print("   %-70s %s" % (a, b))
if out_json:
    json.dump(summary, open(out_json, "w"), indent=1)
