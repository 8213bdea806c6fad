"""Device-resident step of a bench workload, phase times only (A/B timing of kernel variants selected by environment
variables that the library reads when it loads). Usage: python tools/prof_phase.py [workload] [steps]"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench

PKG = bench.PKG
abi, gop, synth, me = (importlib.import_module(PKG + "." + m) for m in ("abi", "gop", "synth", "me"))
if os.environ.get("PROF_LIB"):  # an experimental build of the library (tools/build_variant.sh)
    abi.LIB_PATH = os.path.join(ROOT, os.environ["PROF_LIB"])
import torch

name = sys.argv[1] if len(sys.argv) > 1 else "1080p_p8"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
wl = bench.WORKLOADS[name]
W, H, F = wl["width"], wl["height"], wl["frames"]
seq = synth.synth_sequence(W, H, F, seed=1)
dev = torch.from_numpy(seq).cuda()
ses = me.MeSession(0, W, H, capacity=F)
st = torch.cuda.Stream()
torch.cuda.set_stream(st)
ses.set_stream(st.cuda_stream)
specs = bench.build_specs(abi, gop, wl, ses.derive_params)
if len(sys.argv) > 3:  # only the first N ME pictures (small-batch behaviour of the launch chain)
    specs = specs[:int(sys.argv[3])]
jobs, outs = ses.make_jobs(specs, deep=False)
tot = None
srch = None
ses.set_search_timing(True)
for i in range(steps + 2):
    ses.put_pictures(list(range(F)), [dev.data_ptr() + f * W * H for f in range(F)], W, device=True)
    ses.launch_jobs(jobs)
    ph = ses.last_phase_ms()
    sm = ses.last_search_ms()
    if i >= 2:
        tot = ph if tot is None else {k: tot[k] + v for k, v in ph.items()}
        srch = sm if srch is None else {k: srch[k] + v for k, v in sm.items()}
work = ses.last_phase_work()
print(name, {k: round(v / steps, 3) for k, v in tot.items()}, "sum %.3f" % (sum(tot.values()) / steps),
      "env", {k: v for k, v in os.environ.items() if k.startswith("SVTME_")})
print("   search kernels alone", {k: round(v / steps, 3) for k, v in srch.items() if v > 0})
print("   work G byte-ops", {k: round(v / 1e9, 2) for k, v in work.items() if v})
ses.close()
