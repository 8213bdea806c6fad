"""One resident step of the bench workload, for ncu (GPU box): python tools/prof_step.py [workload] [steps]"""
import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "svt-av1-mod-by-patman_b200"
me = importlib.import_module(PKG + ".me"); gop = importlib.import_module(PKG + ".gop"); abi = importlib.import_module(PKG + ".abi")
synth = importlib.import_module(PKG + ".synth")
import bench
name = sys.argv[1] if len(sys.argv) > 1 else "1080p_p8"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
wl = bench.WORKLOADS[name]
W, H, F = wl["width"], wl["height"], wl["frames"]
dev = torch.from_numpy(synth.synth_sequence(W, H, F, seed=1)).cuda()
ses = me.MeSession(0, W, H, capacity=F)
specs = bench.build_specs(abi, gop, wl, ses.derive_params)
jobs, outs = ses.make_jobs(specs)
for _ in range(steps):
    ses.put_pictures(list(range(F)), [dev.data_ptr() + f * W * H for f in range(F)], W, device=True)
    ses.launch_jobs(jobs)
    ses.sync()
print("phase_ms", {k: round(v, 3) for k, v in ses.last_phase_ms().items()}, "kernel_ms", round(ses.last_kernel_ms(), 3))
ses.close()
