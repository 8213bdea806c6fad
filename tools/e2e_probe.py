"""Where the end-to-end step goes: raw H2D / D2H rates, put alone, host scatter alone (GPU box)."""
import importlib, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "svt-av1-mod-by-patman_b200"
me = importlib.import_module(PKG + ".me"); gop = importlib.import_module(PKG + ".gop"); abi = importlib.import_module(PKG + ".abi")
import bench
W, H, F = 1920, 1080, 60
wl = bench.WORKLOADS["1080p_p8"]
host = torch.randint(0, 255, (F, H, W), dtype=torch.uint8).pin_memory()
dev = torch.empty_like(host, device="cuda")
def t(fn, n=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
print("torch 1D H2D 124MB ms", t(lambda: dev.copy_(host, non_blocking=True)))
out_h = torch.empty(22 << 20, dtype=torch.uint8).pin_memory(); out_d = torch.empty(22 << 20, dtype=torch.uint8, device="cuda")
print("torch 1D D2H 22MB ms", t(lambda: out_h.copy_(out_d, non_blocking=True)))
a = np.empty(22 << 20, dtype=np.uint8); b = out_h.numpy()
t0 = time.perf_counter(); 
for _ in range(5): a[:] = b
print("host memcpy 22MB ms", (time.perf_counter() - t0) / 5 * 1e3)
ses = me.MeSession(0, W, H, capacity=F)
st = torch.cuda.Stream(); torch.cuda.set_stream(st); ses.set_stream(st.cuda_stream)
fr = list(range(F)); ptrs = [host.data_ptr() + f * W * H for f in fr]
print("put_pictures 60 host ms", t(lambda: ses.put_pictures(fr, ptrs, W)))
dptrs = [dev.data_ptr() + f * W * H for f in fr]
print("put_pictures 60 device ms", t(lambda: ses.put_pictures(fr, dptrs, W, device=True)))
specs = bench.build_specs(abi, gop, wl, ses.derive_params)
jobs, outs = ses.make_jobs(specs)
print("launch only ms", t(lambda: ses.launch_jobs(jobs)))
print("run (launch+fetch+scatter) ms", t(lambda: ses.run_jobs(jobs)))
def sc():
    ses.submit_jobs(jobs); ses.collect_jobs(jobs)
print("submit+collect ms", t(sc))
t0 = time.perf_counter(); ses.put_pictures(fr, ptrs, W); t1 = time.perf_counter(); ses.submit_jobs(jobs); t2 = time.perf_counter()
ses.collect_jobs(jobs); t3 = time.perf_counter()
print("host time: put %.3f submit %.3f collect %.3f ms" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3))
