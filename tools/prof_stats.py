"""Picture statistics (svtme_pictures_statistics) of a bench workload's frames: device time per picture against the HBM
roofline. Usage: python tools/prof_stats.py [workload] [steps]. Algorithmic bytes per picture = the luma rows the variance
reads (every other row for BLOCK_MEAN_PREC_SUB) + the sampled bytes of the 1/16 plane + the outputs."""
import ctypes as C
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
import _util as U

PKG = bench.PKG
abi, synth, me = (importlib.import_module(PKG + "." + m) for m in ("abi", "synth", "me"))
import torch

name = sys.argv[1] if len(sys.argv) > 1 else "1080p_p8"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
wl = bench.WORKLOADS[name]
W, H, F = wl["width"], wl["height"], min(wl["frames"], 32)
seq = synth.synth_sequence(W, H, F, seed=1)
dev = torch.from_numpy(seq).cuda()
n_b64 = ((W + 63) // 64) * ((H + 63) // 64)
ses = me.MeSession(0, W, H, capacity=F)
st = torch.cuda.Stream()
torch.cuda.set_stream(st)
ses.set_stream(st.cuda_stream)
ses.put_pictures(list(range(F)), [dev.data_ptr() + f * W * H for f in range(F)], W, device=True)
p = U.stats_params()  # the reference's defaults: sub-sampled means, all 85 block sizes, 4x4 regions, decimation 4
outs = (abi.SvtMeStatsOut * F)()
keep = []
for i in range(F):
    o, arrays = abi.alloc_stats_out(n_b64, p)
    outs[i] = o
    keep.append(arrays)
nums = (C.c_uint64 * F)(*range(F))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ms, kms, dms = [], [], []
for k in range(steps + 2):
    torch.cuda.synchronize()
    e0.record(st)
    assert ses.lib.svtme_pictures_statistics(ses.h, F, nums, C.byref(p), outs) == 0
    e1.record(st)
    torch.cuda.synchronize()
    a, b = C.c_float(), C.c_float()
    assert ses.lib.svtme_last_statistics_ms(ses.h, C.byref(a), C.byref(b)) == 0
    if k >= 2:
        ms.append(e0.elapsed_time(e1))
        kms.append(a.value)
        dms.append(b.value)
# parity of the last call against the CPU checker (first and last picture)
which = U.best_cpu_checker()
for i in (0, F - 1):
    want = U.cpu_picture_statistics(which, U.CpuPyramid(which, seq[i], i), p)
    for k2 in want:
        assert np.array_equal(want[k2], keep[i][k2]), (i, k2)
per_pic = float(np.median(kms)) / F  # the kernels alone; the synchronous call (download + host scatter) is reported beside it
alg = W * H / 2 + (W // 4) * (H // 4) / 16 + n_b64 * 85 * 2 + 16 * 256 * 4
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
hbm = peaks.get("hbm_gbs", 6650.0)
print(json.dumps({"workload": name, "pictures_per_call": F, "ms_per_call": float(np.median(ms)), "kernels_ms_per_call": float(np.median(kms)),
                  "download_ms_per_call": float(np.median(dms)), "kernels_us_per_picture": per_pic * 1e3,
                  "algorithmic_bytes_per_picture": alg, "achieved_gbs": alg / (per_pic * 1e-3) / 1e9, "hbm_peak_gbs": hbm,
                  "frac": alg / (per_pic * 1e-3) / 1e9 / hbm, "roofline": "kernels of one call (4 launches for all pictures), CUDA events inside the library",
                  "parity": "bit-exact against " + which}))
ses.close()
