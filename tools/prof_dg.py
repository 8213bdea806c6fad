"""Dynamic-GOP detector HME (svtme_dg_detector_hme) between consecutive frames of a bench workload: kernel time against the
integer-SAD roofline (the same VABSDIFF4 peak bench.py measures). Usage: python tools/prof_dg.py [workload] [pairs].
Algorithmic work per b64 = clipped window positions x 256 byte-SADs (16x16 block, every row)."""
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

abi, synth, me = (importlib.import_module(bench.PKG + "." + m) for m in ("abi", "synth", "me"))
import torch

name = sys.argv[1] if len(sys.argv) > 1 else "1080p_p8"
pairs = int(sys.argv[2]) if len(sys.argv) > 2 else 8
wl = bench.WORKLOADS[name]
W, H, F = wl["width"], wl["height"], pairs + 1
res = 4 if W * H < 0x535200 else (5 if W * H < 0x140A000 else 6)  # svt_aom_derive_input_resolution (sequence_control_set.c:113-131)
seq = synth.synth_sequence(W, H, F, seed=1)
dev = torch.from_numpy(seq).cuda()
ses = me.MeSession(0, W, H, capacity=F)
ses.put_pictures(list(range(F)), [dev.data_ptr() + f * W * H for f in range(F)], W, device=True)
which = U.best_cpu_checker()
kms = []
for rep in range(3):
    for k in range(pairs):
        m = abi.SvtMeDgMetrics()
        assert ses.lib.svtme_dg_detector_hme(ses.h, k + 1, k, res, C.byref(m)) == 0
        a, b = C.c_float(), C.c_float()
        assert ses.lib.svtme_last_statistics_ms(ses.h, C.byref(a), C.byref(b)) == 0
        if rep:
            kms.append(a.value)
        if rep == 2 and k in (0, pairs - 1):  # parity of this run against the CPU checker
            want = U.cpu_dg_detector_hme(which, U.CpuPyramid(which, seq[k + 1], k + 1), U.CpuPyramid(which, seq[k], k), res)
            assert (int(m.tot_dist), int(m.tot_cplx), int(m.tot_active), int(m.sum_in_vectors)) == want
# work: clipped windows (hme_level_0's clipping, pad 15) of every b64
w16, h16, S = W // 4, H // 4, 128
work = 0
for by in range((H + 63) // 64):
    for bx in range((W + 63) // 64):
        ox, oy, sw, sh = -S // 2, -S // 2, S, S
        x, y = bx * 16, by * 16
        if x + ox < -15:
            ox = -15 - x
        if x + ox > w16 - 1:
            ox -= (x + ox) - (w16 - 1)
        if x + ox + sw > w16:
            sw = max(1, sw - ((x + ox + sw) - w16))
        sw = sw if sw < 8 else sw & ~7
        if y + oy < -15:
            oy = -15 - y
        if y + oy > h16 - 1:
            oy -= (y + oy) - (h16 - 1)
        if y + oy + sh > h16:
            sh = max(1, sh - ((y + oy + sh) - h16))
        work += sw * sh * 256
vsad = float(ses.lib.svtme_measure_vabsdiff4_ginst_per_s())
peak = vsad * 4 / 1e3
ms = float(np.median(kms))
print(json.dumps({"workload": name, "input_resolution": res, "window": S, "kernel_ms_per_pair": ms, "work_byte_ops_per_pair": work,
                  "achieved_tbyte_sad_s": work / (ms * 1e-3) / 1e12, "peak_tbyte_sad_s": peak, "frac": work / (ms * 1e-3) / 1e12 / peak,
                  "parity": "bit-exact against " + which}))
ses.close()
