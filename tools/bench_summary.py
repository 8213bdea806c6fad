"""Prints the numbers of bench.py JSON lines that matter while tuning kernels."""
import json
import sys

for f in sys.argv[1:]:
    try:
        d = json.loads([x for x in open(f).read().splitlines() if x.startswith("{")][-1])
    except Exception as e:  # noqa: BLE001
        print(f, "no JSON line:", e)
        continue
    r = d.get("roofline", {})
    print("== %s: %.3f ms/step resident (%.2f M SB/s), e2e %.3f ms (streamed %.3f), launches %s, parity %s" % (
        f, d["ms_per_step"], d["value"] / 1e6, d["e2e"].get("drained", {}).get("ms_per_step", 0), d["e2e"].get("ms_per_step", 0),
        d.get("gpu_launches"), d.get("parity", {}).get("bit_exact")))
    if r:
        print("   dominant: %s frac %.3f | pipeline frac %.3f (%.3f ms)" % (r["kernel"][:60], r["frac"], r["pipeline"]["frac"], r["pipeline"]["ms"]))
        print("   phase_ms  ", r["phase_ms"])
        print("   phase_frac", r["phase_frac"])
    c = d.get("cpu_baseline", {})
    print("   cpu %.0f SB/s on %s cores (eff %.2f); clocks %s" % (c.get("value", 0), c.get("cores"), c.get("parallel_efficiency", 0), d.get("clocks")))
