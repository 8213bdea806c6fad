"""Host-side timeline of one pipelined end-to-end step (GPU box): when each put/submit/collect returns."""
import importlib, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "svt-av1-mod-by-patman_b200"
me = importlib.import_module(PKG + ".me"); gop = importlib.import_module(PKG + ".gop"); abi = importlib.import_module(PKG + ".abi")
synth = importlib.import_module(PKG + ".synth")
import bench
W, H, F = 1920, 1080, 60
wl = bench.WORKLOADS["1080p_p8"]
sizes = [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "4,12,16").split(",")]
depth = int(sys.argv[2]) if len(sys.argv) > 2 else 2
host = torch.from_numpy(synth.synth_sequence(W, H, F, seed=1)).pin_memory()
ses = me.MeSession(0, W, H, capacity=F)
st = torch.cuda.Stream(); torch.cuda.set_stream(st); ses.set_stream(st.cuda_stream)
specs = bench.build_specs(abi, gop, wl, ses.derive_params)
chunks, pos, have = [], 0, set()
while pos < len(specs):
    n = sizes[min(len(chunks), len(sizes) - 1)]
    part = specs[pos:pos + n]; pos += n
    need = []
    for _, cur, refs in part:
        for f in [cur] + list(refs.values()):
            if f not in have:
                have.add(f); need.append(f)
    cj, co = ses.make_jobs(part)
    chunks.append((need, [host.data_ptr() + f * W * H for f in need], cj, co))
def step(log):
    import ctypes as C
    ses.lib.svtme_debug_mark(ses.h)
    t0 = time.perf_counter()
    def mark(what):
        log.append((what, (time.perf_counter() - t0) * 1e3))
        if what.startswith("collect"):
            v = (C.c_float * 3)()
            ses.lib.svtme_debug_last_span(ses.h, v)
            log.append(("  gpu[%.2f-%.2f d2h %.2f]" % (v[0], v[1], v[2]), (time.perf_counter() - t0) * 1e3))
    for c, (need, ptrs, cj, _) in enumerate(chunks):
        ses.put_pictures(need, ptrs, W); mark("put%d(%d)" % (c, len(need)))
        ses.submit_jobs(cj); mark("submit%d(%d)" % (c, len(cj)))
        if c >= depth:
            ses.collect_jobs(chunks[c - depth][2]); mark("collect%d" % (c - depth))
    for c in range(max(0, len(chunks) - depth), len(chunks)):
        ses.collect_jobs(chunks[c][2]); mark("collect%d" % c)
for _ in range(3):
    step([])
torch.cuda.synchronize()
log = []
step(log)
print(" ".join("%s@%.2f" % x for x in log))
# each chunk alone, device time and wall time
for c, (need, ptrs, cj, _) in enumerate(chunks):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    ses.submit_jobs(cj); ses.collect_jobs(cj)
    print("chunk %d alone: %d jobs, wall %.2f ms, kernels %.2f ms" % (c, len(cj), (time.perf_counter() - t0) * 1e3, ses.last_kernel_ms()))
