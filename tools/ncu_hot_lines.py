#!/usr/bin/env python3
"""Per-source-line and per-opcode summary of an `ncu --set full --import-source on` report.

ncu's source page (CSV) lists SASS instructions with their sampling counters but not the CUDA
line they came from; nvdisasm --print-line-info on the cubin of the same build gives the
offset -> file:line map. This joins the two.

  python tools/ncu_hot_lines.py gpurun_out/prof.ncu-rep svt-av1-mod-by-patman_b200/csrc/libsvtme_b200.so \
         me_stage_kernel [--top 40] [--json out.json]
"""
import argparse, collections, csv, io, json, os, re, subprocess, sys, tempfile


def sh(cmd, **kw):
    return subprocess.run(cmd, check=True, capture_output=True, text=True, **kw).stdout


def line_map(so_path, kernel):
    """{function name: {sass offset: 'file:line'}} for every function in the cubin(s) of the .so."""
    out = {}
    with tempfile.TemporaryDirectory() as td:
        sh(["cuobjdump", "-xelf", "all", os.path.abspath(so_path)], cwd=td)
        for f in sorted(os.listdir(td)):
            if not f.endswith(".cubin"):
                continue
            try:
                txt = sh(["nvdisasm", "--print-line-info", "-c", os.path.join(td, f)])
            except subprocess.CalledProcessError:
                continue
            if kernel not in txt:
                continue
            fn, cur = None, None
            for ln in txt.splitlines():
                m = re.match(r"\s*\.text\.(\S+):", ln)
                if m:
                    fn = m.group(1)
                    out.setdefault(fn, {})
                    cur = None
                    continue
                m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
                if m:
                    if "inlined at" in ln:  # keep the innermost user line
                        pass
                    cur = "%s:%s" % (os.path.basename(m.group(1)), m.group(2))
                    continue
                m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
                if m and fn:
                    out[fn][int(m.group(1), 16)] = (cur, m.group(2))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("so")
    ap.add_argument("kernel")
    ap.add_argument("--top", type=int, default=40)
    ap.add_argument("--json")
    a = ap.parse_args()

    src = sh(["ncu", "-i", a.rep, "--page", "source", "--csv", "--print-source", "sass"])
    # the CSV may hold several functions (device functions called by the kernel): ncu prints one
    # header per function preceded by a name line
    rows, header, fn = [], None, None
    for rec in csv.reader(io.StringIO(src)):
        if not rec:
            continue
        if rec[0] in ("Address", "#") or "Source" in rec:
            header = rec
            continue
        if header is None or len(rec) != len(header):
            # function title line
            fn = " ".join(rec)
            continue
        rows.append((fn, dict(zip(header, rec))))
    if not rows:
        sys.exit("no source rows in report")
    keys = rows[0][1].keys()
    k_samp = next((k for k in keys if k.strip() == "# Samples" or k.startswith("Warp Stall Sampling (All")), None)
    k_exec = next((k for k in keys if k.startswith("# Warp Instructions Executed") or k == "Instructions Executed"), None)
    k_src = "Source"
    lm = line_map(a.so, a.kernel)
    kmap = next((mp for name, mp in lm.items() if a.kernel in name), {})

    def parse_num(s):
        try:
            return float(s.replace(",", ""))
        except Exception:
            return 0.0

    by_line = collections.Counter()
    by_line_exec = collections.Counter()
    by_op = collections.Counter()
    by_op_samp = collections.Counter()
    tot_s = tot_e = 0.0
    # address -> offset: ncu prints absolute addresses; offsets restart per function
    base = {}
    for fn, r in rows:
        addr = int(r["Address"], 16) if r.get("Address", "").startswith("0x") or re.match(r"^[0-9a-f]+$", r.get("Address", "")) else None
        if addr is None:
            continue
        base.setdefault(fn, addr)
    for fn, r in rows:
        try:
            addr = int(r["Address"], 16)
        except Exception:
            continue
        off = addr - base[fn]
        smp, ex = parse_num(r.get(k_samp, "0")), parse_num(r.get(k_exec, "0"))
        tot_s += smp
        tot_e += ex
        sass = r[k_src].strip()
        op = sass.split()[0] if sass else "?"
        if op.startswith("@"):
            op = sass.split()[1] if len(sass.split()) > 1 else op
        by_op[op] += ex
        by_op_samp[op] += smp
        hit = kmap.get(off)
        line = hit[0] if hit else None
        by_line[line] += smp
        by_line_exec[line] += ex

    print("total samples %.0f, warp instructions %.4g" % (tot_s, tot_e))
    print("\n== hot source lines (share of stall samples | share of executed warp instructions)")
    out_lines = []
    for line, smp in by_line.most_common(a.top):
        e = by_line_exec[line]
        print("%6.2f%%  %6.2f%%  %s" % (100 * smp / max(tot_s, 1), 100 * e / max(tot_e, 1), line))
        out_lines.append({"line": line, "samples_pct": 100 * smp / max(tot_s, 1), "inst_pct": 100 * e / max(tot_e, 1)})
    print("\n== opcodes (share of executed warp instructions | share of samples)")
    out_ops = []
    for op, e in by_op.most_common(25):
        print("%6.2f%%  %6.2f%%  %s" % (100 * e / max(tot_e, 1), 100 * by_op_samp[op] / max(tot_s, 1), op))
        out_ops.append({"op": op, "inst_pct": 100 * e / max(tot_e, 1), "samples_pct": 100 * by_op_samp[op] / max(tot_s, 1)})
    if a.json:
        json.dump({"report": os.path.basename(a.rep), "kernel": a.kernel, "total_samples": tot_s,
                   "warp_instructions": tot_e, "lines": out_lines, "opcodes": out_ops}, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
