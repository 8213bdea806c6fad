#!/usr/bin/env python
"""bench.py — open-loop ME throughput (SBs/sec, frames/sec) of the B200 path, the driver contract.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload 1080p_p8|4k_p4] [--impl reference]

One "step" = one pass of the hot path over one batch of synthetic input: the pre-ME pyramid of all
frames of the sequence + the open-loop ME (pre-HME, HME L0-L2, full-pel 85-PU search, candidate
arrays) of every non-intra picture against its references.
  value  : SBs/s with the raw luma frames already resident in HBM (device-timed, CUDA events)
  e2e    : same metric through the public C ABI with HOST buffers: per step every frame is copied
           H2D from pinned memory and every result table is copied back D2H, inside the timed region
  N > 1  : one process per GPU; rank r owns its own independent stream (weak scaling, no data-path
           collective); time = max over ranks.
--impl reference times the reference's own CPU implementation of the same path (oracle/_ref: the
unmodified reference sources with its AVX2 kernels; the oracle port if that library is absent) on
all host threads, on a bounded sample of the same workload.
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# One process per GPU shares the host cores: the library's result scatter gets cores / (2 x ranks) threads (1..4) and
# idle OpenMP workers sleep instead of spinning (both read when the libraries load, i.e. before `import torch`).
_world = int(os.environ.get("WORLD_SIZE", "1"))
os.environ.setdefault("SVTME_SCATTER_THREADS", str(max(1, min(8, (os.cpu_count() or 1) // (2 * _world)))))
os.environ.setdefault("OMP_WAIT_POLICY", "passive")
PKG = "svt-av1-mod-by-patman_b200"

WORKLOADS = {
    # configs[1] of BASELINE.json: 1080p 8-bit, preset 8, 60 frames, refs 3+2 (mrp 9), 5 hierarchical levels
    "1080p_p8": dict(width=1920, height=1080, preset=8, frames=60, levels=5, n_l0=3, n_l1=2),
    # configs[2]: 4K (the ME stage sees the 8 MSBs of 10-bit input), preset 4, 120 frames, refs 4+3
    # source depth 10: 16-bit samples = the 8-bit pattern << 2 plus two random LSBs; the boundary takes them as they are
    # (svtme_put_pictures_10bit) and the planes hold sample >> 2, like the reference's 8-bit ME buffer
    "4k_p4": dict(width=3840, height=2160, preset=4, frames=120, levels=5, n_l0=4, n_l1=3, bits=10),
    # configs[3]: 8K 10-bit, preset 6, mini-GOP 16 (4 hierarchical levels), refs 3+2; meant for --sharding picture on 8 GPUs
    "8k_p6": dict(width=7680, height=4320, preset=6, frames=33, levels=4, n_l0=3, n_l1=2, bits=10),
    "small": dict(width=640, height=360, preset=8, frames=12, levels=3, n_l0=2, n_l1=1),
    # configs[0]: kernel level — full-pel search (open_loop_me_fullpel_search_sblock) of every 64x64 block of a random
    # 1920x1080 frame pair, centre (0,0), SUB_SAD, search areas of test/SadTest.cc:507-518 (handled by run_kernel_workload)
    "kernel_1080p": dict(width=1920, height=1080, kernel=True),
}


# search kernel of each launch slot of the stage pipeline (svtme_stage.cu) for SUB_SAD_SEARCH presets with small
# HME refinement windows, i.e. every workload of this file
SLOT_KERNELS = {"zz_sad": "points_kernel", "prehme": "strip_kernel<0>", "hme_l0": "strip_kernel<1>", "hme_l1": "refine_kernel<8>",
                "hme_l2_centre": "refine_kernel<16>", "integer_search": "int_kernel<true> (+ int_kernel<false> for parked areas)",
                "finish": "finish_kernel"}


def build_specs(abi, gop, wl, derive_fn):
    """(params, cur, refs) per non-intra picture, decode order."""
    U = importlib.import_module(PKG + ".me")
    specs = []
    for j in gop.ra_jobs(wl["frames"], wl["levels"], wl["n_l0"], wl["n_l1"]):
        d = U.derive_input(wl["preset"], wl["width"], wl["height"], hierarchical_levels=wl["levels"],
                           temporal_layer_index=j.temporal_layer, is_ref=int(j.is_ref), ref_list0_count_try=j.n_l0,
                           ref_list1_count_try=j.n_l1, ref_count_used_list0=wl["n_l0"], ref_count_used_list1=wl["n_l1"])
        p = derive_fn(d)
        U.finish_params(p, d, j.picture_number, j.refs)
        specs.append((p, j.picture_number, j.refs))
    return specs


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons sampled DURING the timed region (NVML in-process, ~2 ms per sample;
    nvidia-smi subprocess as the fallback)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            self.dev = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_sm = int(pynvml.nvmlDeviceGetMaxClockInfo(self.dev, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nvml = None

    def run(self):
        if self.nvml is not None:
            n = self.nvml
            bits = [("hw_slowdown", n.nvmlClocksThrottleReasonHwSlowdown), ("hw_thermal_slowdown", n.nvmlClocksThrottleReasonHwThermalSlowdown),
                    ("sw_thermal_slowdown", n.nvmlClocksThrottleReasonSwThermalSlowdown), ("sw_power_cap", n.nvmlClocksThrottleReasonSwPowerCap)]
            while not self.stop_flag:
                try:
                    sm = int(n.nvmlDeviceGetClockInfo(self.dev, n.NVML_CLOCK_SM))
                    r = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.dev))
                    self.rows.append([str(sm), str(self.max_sm)] + ["Active" if r & b else "Not Active" for _, b in bits])
                except Exception:
                    pass
                time.sleep(0.01)
            return
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": int(self.rows[0][1]) if self.rows[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(self.rows), "source": "nvml" if self.nvml is not None else "nvidia-smi"}


def cpu_reference_run(wl, specs, seq, n_pictures, threads, min_seconds=10.0):
    """ME of the first n_pictures jobs (+ the pyramids of the frames they touch) on the host: the compiled reference with
    its AVX2 kernels if oracle/_ref is present, else the oracle port. Threading is native (oracle/mt_pool.h): a pthread
    pool pulls (picture, b64-row) tasks from one queue — the reference's own ME-thread / segment scheme — and builds
    the pyramids one picture per task. Returns throughput, the 1-thread time of the same sample (parallel efficiency)
    and the result arrays of the sample (for the bit-exactness check against the GPU outputs)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _util as U  # oracle drivers: CPU-baseline leg only
    which = U.best_cpu_checker()
    kind = "reference" if which == "ref" else "port"
    simd = "avx2" if (which == "ref" and U.ref_lib().ref_simd_level() == 1) else "c"
    jobs = specs[:n_pictures]
    need = sorted({n for _, cur, refs in jobs for n in [cur] + list(refs.values())})
    frames = {n: seq[n] for n in need}
    n_b64 = ((wl["width"] + 63) // 64) * ((wl["height"] + 63) // 64)
    pyrs = U.CpuPyramidSet(which).build(frames, threads)  # allocates the planes (untimed, like the encoder's buffer pools)
    cj, outs = U.cpu_mt_jobs(which, jobs, pyrs, n_b64)

    def one_pass(nthreads):
        t0 = time.perf_counter()
        pyrs.build(frames, nthreads)
        t1 = time.perf_counter()
        U.cpu_me_pictures_mt(which, cj, nthreads)
        return t1 - t0, time.perf_counter() - t1

    one_pass(threads)  # warm-up
    t_pyr = t_me = 0.0
    reps = 0
    t_begin = time.perf_counter()
    while True:
        a, b = one_pass(threads)
        t_pyr, t_me, reps = t_pyr + a, t_me + b, reps + 1
        if time.perf_counter() - t_begin >= min_seconds:
            break
    t_pyr, t_me = t_pyr / reps, t_me / reps
    p1, m1 = one_pass(1) if threads > 1 else (t_pyr, t_me)
    one_pass(threads)  # leave the multi-threaded results in `outs` (identical anyway)
    # pyramid cost is charged per ME'd picture like on the GPU side (one pyramid per frame of the sequence)
    per_pic = t_me / len(jobs) + t_pyr / len(need)
    per_pic_1 = m1 / len(jobs) + p1 / len(need)
    return dict(sbs_per_s=n_b64 / per_pic, fps=1.0 / per_pic, kind=kind, simd=simd, seconds=time.perf_counter() - t_begin,
                pictures=len(jobs), pyramids=len(need), reps=reps, threads=threads, sbs_per_s_1thread=n_b64 / per_pic_1,
                parallel_efficiency=per_pic_1 / (threads * per_pic), outs=outs, keep=(pyrs, cj))


def cpu_sample_text(cpu):
    return ("first %d ME pictures of the workload + the %d pyramids they touch, %d passes in %.1f s, %s kernels, %d native "
            "threads pulling (picture, b64-row) tasks from one queue; 1 thread: %.0f SB/s -> parallel efficiency %.2f"
            % (cpu["pictures"], cpu["pyramids"], cpu["reps"], cpu["seconds"], cpu["simd"], cpu["threads"], cpu["sbs_per_s_1thread"],
               cpu["parallel_efficiency"]))


def compare_outputs(cpu_outs, gpu_outs_list):
    """Bit-exactness of the GPU result arrays against the CPU sample. gpu_outs_list: [(label, [arrays per picture])]."""
    arrays, bad = 0, []
    for label, gouts in gpu_outs_list:
        for i, want in enumerate(cpu_outs):
            for k, w in want.items():
                arrays += 1
                if not np.array_equal(w, gouts[i][k]):
                    bad.append("%s picture#%d %s" % (label, i, k))
    return arrays, bad


KERNEL_WINDOWS = [(16, 5), (64, 32), (128, 64), (256, 256)]


def kernel_items(abi, W, H, stride, org_x, org_y, sa_w0, sa_h0):
    """One SvtMeFullpelItem per 64x64 block: search area centred on MV (0,0) and clipped like integer_search_b64 clips it
    (motion_estimation.c:1470-1561: at most 63 pixels into the padding, block origin inside the picture, width a multiple
    of 8 once it is 8 or more). Returns (ctypes array, [(ox, oy, w, h)], positions in total)."""
    cols, rows = (W + 63) // 64, (H + 63) // 64
    items = (abi.SvtMeFullpelItem * (cols * rows))()
    geo, positions = [], 0
    for by in range(rows):
        for bx in range(cols):
            x0, y0, w, h = bx * 64, by * 64, sa_w0, sa_h0
            ox, oy = -(w >> 1), -(h >> 1)
            if x0 + ox < -63:
                ox = -63 - x0
            if x0 + ox > W - 1:
                ox -= (x0 + ox) - (W - 1)
            if x0 + ox + w > W:
                w = max(1, w - ((x0 + ox + w) - W))
            w = w if w < 8 else w & ~7
            if y0 + oy < -63:
                oy = -63 - y0
            if y0 + oy > H - 1:
                oy -= (y0 + oy) - (H - 1)
            if y0 + oy + h > H:
                h = max(1, h - ((y0 + oy + h) - H))
            it = items[by * cols + bx]
            it.src_offset = (org_y + y0) * stride + org_x + x0
            it.ref_offset = (org_y + y0 + oy) * stride + org_x + x0 + ox
            it.x_search_area_origin, it.y_search_area_origin, it.search_area_width, it.search_area_height = ox, oy, w, h
            geo.append((ox, oy, w, h))
            positions += w * h
    return items, geo, positions


def run_kernel_workload(args, wl):
    """BASELINE configs[0]: the full-pel engine alone, through the batched kernel-level entry point of the C ABI."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _util as U  # oracle drivers: parity check and CPU-baseline leg only
    abi = importlib.import_module(PKG + ".abi")
    me = importlib.import_module(PKG + ".me")
    W, H = wl["width"], wl["height"]
    n_b64 = ((W + 63) // 64) * ((H + 63) // 64)
    rng = np.random.default_rng(13596)
    frames = rng.integers(0, 256, size=(2, H, W), dtype=np.uint8)  # i.i.d. uniform pixels (SadTest RANDOM pattern)
    host = torch.from_numpy(frames).pin_memory()
    ses = me.MeSession(0, W, H, capacity=2)
    lib = ses.lib
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    ses.set_stream(stream.cuda_stream)
    ses.put_pictures([0, 1], [host.data_ptr(), host.data_ptr() + W * H], W)
    ses.sync()
    cur_ptr, _ = ses.plane_device_ptr(0, 0)
    ref_ptr, _ = ses.plane_device_ptr(1, 0)
    org_x, org_y, stride = 128, 72, (W + 256 + 127) & ~127  # level-0 geometry of the resident planes
    vsad = float(lib.svtme_measure_vabsdiff4_ginst_per_s())
    int_peak = vsad * 4 / 1e3
    pyr = [U.CpuPyramid("orc", frames[k], k) for k in range(2)]
    orc = U.oracle_lib()
    results, parity_blocks, parity_bad = {}, 0, []
    sampler = ClockSampler(0)
    sampler.start()
    for (sw, sh) in KERNEL_WINDOWS:
        items, geo, positions = kernel_items(abi, W, H, stride, org_x, org_y, sw, sh)
        h_items = torch.frombuffer(bytearray(bytes(items)), dtype=torch.uint8).pin_memory()
        d_items = h_items.cuda()
        init_sad = np.full((n_b64, 85), abi.MAX_SAD_VALUE, dtype=np.uint32)
        h_sad = torch.from_numpy(init_sad.view(np.int32).copy()).pin_memory()
        h_mv = torch.zeros((n_b64, 85), dtype=torch.int32).pin_memory()
        d_sad, d_mv = h_sad.cuda(), h_mv.cuda()
        out_sad, out_mv = torch.empty_like(h_sad).pin_memory(), torch.empty_like(h_mv).pin_memory()

        def launch():
            rc = lib.svt_open_loop_me_fullpel_search_sblock_b200_batch(stream.cuda_stream, n_b64, cur_ptr, stride, ref_ptr, stride,
                                                                       d_items.data_ptr(), 1, d_sad.data_ptr(), d_mv.data_ptr())
            assert rc == 0, lib.svtme_last_error()

        def e2e_step():
            # host buffers in, host tables out: both frames, the items and the initial tables go up, the tables come back
            ses.put_pictures([0, 1], [host.data_ptr(), host.data_ptr() + W * H], W)
            ses.sync()
            d_items.copy_(h_items, non_blocking=True)
            d_sad.copy_(h_sad, non_blocking=True)
            d_mv.copy_(h_mv, non_blocking=True)
            launch()
            out_sad.copy_(d_sad, non_blocking=True)
            out_mv.copy_(d_mv, non_blocking=True)
            stream.synchronize()

        for _ in range(max(3, args.warmup)):
            launch()
        torch.cuda.synchronize()
        reps = max(args.steps, 1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            launch()
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        e2e_step()
        t0 = time.perf_counter()
        for _ in range(reps):
            e2e_step()
        ms_e2e = (time.perf_counter() - t0) * 1e3 / reps
        work = positions * 2048
        got_sad, got_mv = out_sad.numpy().view(np.uint32), out_mv.numpy().view(np.uint32)
        # parity: a sample of blocks against the oracle restatement of open_loop_me_fullpel_search_sblock
        step_b = max(1, n_b64 // (6 if sw * sh > 10000 else 40))
        full = pyr[1].pyr.full
        for b in range(0, n_b64, step_b):
            ox, oy, w, h = geo[b]
            bx, by = b % ((W + 63) // 64), b // ((W + 63) // 64)
            bs = np.full(85, abi.MAX_SAD_VALUE, dtype=np.uint32)
            bm = np.zeros(85, dtype=np.uint32)
            cf = pyr[0].pyr.full
            src = cf.buffer_y + (cf.org_y + 64 * by) * cf.stride_y + cf.org_x + 64 * bx
            ref = full.buffer_y + (full.org_y + 64 * by + oy) * full.stride_y + full.org_x + 64 * bx + ox
            orc.orc_fullpel_search_sblock(C.c_void_p(src), C.c_uint32(cf.stride_y), C.c_void_p(ref), C.c_uint32(full.stride_y), C.c_int16(ox),
                                          C.c_int16(oy), C.c_uint32(w), C.c_uint32(h), C.c_uint8(1), bs.ctypes.data_as(C.c_void_p),
                                          bm.ctypes.data_as(C.c_void_p))
            parity_blocks += 1
            if not (np.array_equal(bs, got_sad[b]) and np.array_equal(bm, got_mv[b])):
                parity_bad.append("%dx%d block %d" % (sw, sh, b))
        results["%dx%d" % (sw, sh)] = {"ms_per_launch": ms, "sbs_per_s": n_b64 / (ms * 1e-3), "positions": positions,
                                       "work_byte_ops": work, "achieved_tbyte_sad_per_s": work / (ms * 1e-3) / 1e12,
                                       "frac": work / (ms * 1e-3) / 1e12 / int_peak, "e2e_ms": ms_e2e,
                                       "e2e_sbs_per_s": n_b64 / (ms_e2e * 1e-3)}
    sampler.stop_flag = True
    sampler.join(timeout=2)
    # CPU baseline: the oracle port of the same search on all host cores, 64x32 areas, a bounded sample of blocks
    threads = os.cpu_count() or 1
    items, geo, positions = kernel_items(abi, W, H, stride, org_x, org_y, 64, 32)
    cf, rf = pyr[0].pyr.full, pyr[1].pyr.full

    def cpu_blocks(blocks):
        bs, bm = np.empty(85, dtype=np.uint32), np.empty(85, dtype=np.uint32)
        for b in blocks:
            ox, oy, w, h = geo[b]
            bx, by = b % ((W + 63) // 64), b // ((W + 63) // 64)
            bs[:] = abi.MAX_SAD_VALUE
            bm[:] = 0
            orc.orc_fullpel_search_sblock(C.c_void_p(cf.buffer_y + (cf.org_y + 64 * by) * cf.stride_y + cf.org_x + 64 * bx), C.c_uint32(cf.stride_y),
                                          C.c_void_p(rf.buffer_y + (rf.org_y + 64 * by + oy) * rf.stride_y + rf.org_x + 64 * bx + ox),
                                          C.c_uint32(rf.stride_y), C.c_int16(ox), C.c_int16(oy), C.c_uint32(w), C.c_uint32(h), C.c_uint8(1),
                                          bs.ctypes.data_as(C.c_void_p), bm.ctypes.data_as(C.c_void_p))

    t0, done = time.perf_counter(), 0
    while time.perf_counter() - t0 < args.cpu_seconds:
        ths = [threading.Thread(target=cpu_blocks, args=(range(t, n_b64, threads),)) for t in range(threads)]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        done += n_b64
    cpu_sbs = done / (time.perf_counter() - t0)
    head = results["64x32"]
    line = {"metric": "open-loop ME SBs/sec", "value": head["sbs_per_s"], "unit": "SB/s", "n_gpus": 1, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": head["ms_per_launch"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "kernel_1080p (BASELINE configs[0]): full-pel search of all %d 64x64 blocks of a random 1920x1080 "
                                   "frame pair (i.i.d. uniform pixels), centre (0,0), SUB_SAD, 85 PUs; headline = 64x32 search areas, "
                                   "the other areas of test/SadTest.cc under 'windows'" % n_b64,
                       "l2": "the two padded planes (4.7 MB) stay in L2: this workload measures the SAD engine, not HBM"},
            "e2e": {"value": head["e2e_sbs_per_s"], "unit": "SB/s", "h2d_bytes_per_step": 2 * W * H + n_b64 * (16 + 2 * 340),
                    "d2h_bytes_per_step": n_b64 * 2 * 340, "ms_per_step": head["e2e_ms"],
                    "api": "svtme_put_pictures (both frames, pinned host) + svt_open_loop_me_fullpel_search_sblock_b200_batch + download "
                           "of the 85-entry tables"},
            "gpu_launches": int(me.launch_count()),
            "roofline": {"bound": "int_alu", "kernel": "fullpel_batch_kernel (fullpel_search, 64x32 areas)", "achieved": head["achieved_tbyte_sad_per_s"],
                         "peak": int_peak, "unit": "Tbyte-SAD/s", "frac": head["frac"], "traffic": None,
                         "peak_source": "VABSDIFF4.U8.ACC issue rate measured live in this run (%.0f Ginst/s x 4 bytes)" % vsad},
            "windows": results,
            "cpu_baseline": {"value": cpu_sbs, "unit": "SB/s", "cores": threads, "kind": "port",
                             "sample": "orc_fullpel_search_sblock (oracle restatement, plain C) on all %d blocks with 64x32 areas, %d passes, "
                                       "%d threads" % (n_b64, done // n_b64, threads)},
            "parity": {"blocks": parity_blocks, "bit_exact": not parity_bad, "against": "oracle port (orc_fullpel_search_sblock)",
                       **({"mismatches": parity_bad[:8]} if parity_bad else {})},
            "clocks": sampler.summary()}
    print(json.dumps(line))
    ses.close()
    return 1 if parity_bad else 0


def main():
    if os.environ.get("SVTME_BENCH_LIB"):  # development aid: an experimental build of the library (tools/build_variant.sh)
        importlib.import_module(PKG + ".abi").LIB_PATH = os.path.join(ROOT, os.environ["SVTME_BENCH_LIB"])
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="1080p_p8", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--sharding", default="stream", choices=["stream", "picture"],
                    help="N>1: 'stream' = one independent stream per rank (weak scaling, BASELINE configs[4]); "
                         "'picture' = ONE stream, pictures dealt round-robin to the ranks (strong scaling, configs[3])")
    ap.add_argument("--pyramids", default="rebuild", choices=["rebuild", "nccl"],
                    help="--sharding picture: 'rebuild' = every GPU builds the pyramids it needs from the frames; 'nccl' = "
                         "picture p is uploaded and built on GPU p %% N only and sent (NCCL over NVLink) "
                         "into the ring slots of the GPUs that reference it")
    ap.add_argument("--e2e-chunks", default="32", help="pictures per chunk (batch) of the pipelined end-to-end run "
                    "(decode order; the last value repeats): one mini-GOP of the look-ahead window per batch")
    ap.add_argument("--e2e-chunks-shared", default="4,8,16", help="--pyramids nccl: pictures PER RANK in each global chunk of the "
                    "pipelined end-to-end run (the last value repeats)")
    ap.add_argument("--e2e-source", default="msb8", choices=["msb8", "samples16"],
                    help="10-bit workloads, end-to-end run: 'msb8' uploads the 8-bit MSB plane the encoder already holds on the host "
                         "(input_padded_pic->buffer_y, filled by svt_aom_un_pack2d in resource coordination, "
                         "resource_coordination_process.c:857,1022) — what the reference's ME reads; 'samples16' uploads the "
                         "16-bit samples and lets the pyramid kernel take the 8 MSBs")
    ap.add_argument("--pin-results", type=int, default=1, help="end-to-end run: 1 = the callers' three big result tables per picture "
                    "are page-locked (svtme_pin_host_memory) and filled by the download itself; 0 = pageable arrays, filled from "
                    "the library's pinned mirror by svtme_collect_results")
    ap.add_argument("--timeline", action="store_true", help="rank 0: after the timed runs, one more drained end-to-end step with host "
                    "timestamps of every put / exchange / submit / collect and the device spans of every batch (stderr)")
    ap.add_argument("--cpu-pictures", type=int, default=0, help="pictures in the bounded CPU sample (0: 16 at 1080p, 8 at 4K, 4 at 8K)")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="repeat the CPU sample for about this long")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if wl.get("kernel"):
        if args.impl == "reference" or int(os.environ.get("RANK", "0")) != 0:
            if args.impl == "reference" and int(os.environ.get("RANK", "0")) == 0:
                print(json.dumps({"impl": "reference", "unavailable": "the kernel-level workload times its CPU port inside the b200 arm (cpu_baseline)"}))
            return 0
        return run_kernel_workload(args, wl)
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    W, H, F = wl["width"], wl["height"], wl["frames"]
    n_b64 = ((W + 63) // 64) * ((H + 63) // 64)
    depth_txt = "10-bit source (16-bit samples, ME planes = sample >> 2)" if wl.get("bits") == 10 else "8-bit luma"
    config = {"workload": "%s: %dx%d %s, preset %d, %d frames, refs %d+%d, %d hierarchical levels, synthetic "
                          "translating texture + moving rectangles + noise" % (args.workload, W, H, depth_txt, wl["preset"], F,
                                                                             wl["n_l0"], wl["n_l1"], wl["levels"]),
              "b64_per_frame": n_b64, "me_pictures_per_step": F - 1, "pyramids_per_step": F,
              "l2": "inputs larger than L2: each step touches %.0f MB of frames+pyramids (L2 126 MB)" % (F * W * H * 2.7 / 1e6),
              "sharding": "one independent stream per GPU (weak)" if args.sharding == "stream" else
                          "one stream, pictures round-robin over GPUs in decode order, " +
                          ("each GPU builds the pyramids it needs (strong)" if args.pyramids == "rebuild" else
                           "picture p built on GPU p % N and sent over NCCL/NVLink to the GPUs that reference it (strong)")}
    abi = importlib.import_module(PKG + ".abi")
    gop = importlib.import_module(PKG + ".gop")
    synth = importlib.import_module(PKG + ".synth")
    threads = os.cpu_count() or 1
    cpu_pictures = args.cpu_pictures or (16 if W * H <= 1920 * 1088 else (8 if W * H <= 3840 * 2160 else 4))

    if args.impl == "reference":
        if rank != 0:
            return 0
        # parameters from the reference's own svt_aom_sig_deriv_me (oracle/_ref); only the oracle-port fallback (no compiled
        # reference on this machine) has to borrow the product's host-side derivation, which loads the product library
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import _util as U
        if U.best_cpu_checker() == "ref":
            rl = U.ref_lib()

            def derive(d):
                p = abi.SvtMeParams()
                assert rl.ref_derive_params(C.byref(d), C.byref(p)) == 0
                return p
        else:
            lib = abi.load_library()

            def derive(d):
                p = abi.SvtMeParams()
                assert lib.svtme_derive_params(C.byref(d), C.byref(p)) == 0
                return p
        specs = build_specs(abi, gop, wl, derive)
        seq = synth.synth_sequence(W, H, F, seed=1)
        res = None
        times = []
        for i in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            res = cpu_reference_run(wl, specs, seq, cpu_pictures, threads, args.cpu_seconds)
            if i >= args.warmup:
                times.append((time.perf_counter() - t0, res["sbs_per_s"]))
        val = float(np.mean([v for _, v in times])) if times else res["sbs_per_s"]
        line = {"impl": "reference", "metric": "open-loop ME SBs/sec", "value": val, "unit": "SB/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean([t for t, _ in times])) if times else None,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": config,
                "frames_per_s": val / n_b64,
                "cpu_baseline": {"value": val, "unit": "SB/s", "cores": threads, "kind": res["kind"],
                                 "parallel_efficiency": res["parallel_efficiency"], "value_1thread": res["sbs_per_s_1thread"],
                                 "sample": cpu_sample_text(res)},
                "e2e": {"value": val, "unit": "SB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return 0

    if world > 1 and hasattr(os, "sched_setaffinity") and not os.environ.get("SVTME_NO_AFFINITY"):
        # one process per GPU: give each rank its own slice of the host cores BEFORE it allocates pinned memory (first touch
        # puts the pages on that slice's NUMA node; GPUs are numbered along the sockets on HGX boards), so that uploads do
        # not all cross the socket interconnect
        ncpu = os.cpu_count() or 1
        per = max(1, ncpu // world)
        try:
            os.sched_setaffinity(0, set(range(local_rank * per, min(ncpu, (local_rank + 1) * per))))
        except OSError:
            pass
    import torch
    import torch.distributed as dist
    exit_code = 0
    torch.cuda.set_device(local_rank)
    if world > 1:
        # the pyramid exchange runs while ME kernels with hundreds of thousands of CTAs are resident: its NCCL kernels must
        # not queue behind them (high-priority NCCL streams; the exchange stream below is high priority too)
        os.environ.setdefault("TORCH_NCCL_HIGH_PRIORITY", "1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
    me = importlib.import_module(PKG + ".me")
    shard = importlib.import_module(PKG + ".shard")
    by_picture = args.sharding == "picture" and world > 1
    seq = synth.synth_sequence(W, H, F, seed=1 if by_picture else 1 + rank)
    bits = wl.get("bits", 8)
    if bits == 10:
        lsb = np.random.default_rng(100 + rank).integers(0, 4, size=seq.shape, dtype=np.uint16)
        host = torch.from_numpy((seq.astype(np.uint16) << 2) | lsb).pin_memory()
        del lsb
    else:
        host = torch.from_numpy(seq).pin_memory()
    dev = host.to("cuda", non_blocking=False)
    e2e_bits = 8 if (bits == 10 and args.e2e_source == "msb8") else bits
    host_e2e = torch.from_numpy(seq).pin_memory() if (bits == 10 and e2e_bits == 8) else host
    e2e_frame_bytes = W * H * (2 if e2e_bits == 10 else 1)
    n_slots = -(-F // world) * world if (by_picture and args.pyramids == "nccl") else F  # whole groups of N slots
    # streamed end-to-end run: consecutive steps use alternating sets of ring slots (picture numbers offset by ALT), like a
    # real stream whose new pictures get new pa-reference objects: the next step's uploads do not wait for the launches
    # that still read the previous step's pictures
    ALT = 1 << 20
    ses = me.MeSession(local_rank, W, H, capacity=2 * n_slots if (by_picture and args.pyramids == "nccl") else 2 * F)
    stream = torch.cuda.Stream()  # a non-default stream: the library treats NULL as "use the private stream"
    torch.cuda.set_stream(stream)
    ses.set_stream(stream.cuda_stream)
    specs = build_specs(abi, gop, wl, ses.derive_params)
    share = None
    all_specs = specs
    if by_picture:
        mine = shard.rank_plan(len(specs), world, rank)
        frames_here = shard.pictures_for(specs, mine)
        if args.pyramids == "nccl":
            share = dict(owned=[p for p in range(F) if shard.pyramid_owner(p, world) == rank])
        specs = [specs[i] for i in mine]
    else:
        frames_here = list(range(F))
    jobs, outs = ses.make_jobs(specs, deep=False)
    frame_bytes = W * H * (2 if bits == 10 else 1)
    d2h_bytes = sum(a.nbytes for o in outs for a in o.values())

    if share:
        # picture p -> ring slot p on every rank (slots are handed out in order on a fresh handle), built by rank p % N
        xs = torch.cuda.Stream(priority=-1)  # the exchange stream: NCCL work is ordered against it, never against the host
        # (two sets: pictures p and ALT + p, the second one in slots n_slots + p, for consecutive steps of the streamed run)
        ses.external_write_begin(list(range(n_slots)) + [ALT + p for p in range(n_slots)], xs.cuda_stream)
        pool_ptr = ses.plane_device_ptr(0, 0)[0]
        slot_bytes = sum(ses.plane_device_ptr(0, l)[1] for l in range(3))
        for pnum, slot in ((1, 1), (n_slots - 1, n_slots - 1), (ALT, n_slots), (ALT + n_slots - 1, 2 * n_slots - 1)):
            if ses.plane_device_ptr(pnum, 0)[0] != pool_ptr + slot * slot_bytes:
                raise RuntimeError("ring slots are expected in picture order")
        share.update(xs=xs, pool=pool_ptr, slot_bytes=slot_bytes, cache={}, moved=0)

    def exchange(groups, pictures):
        """All-gather the slots of `groups` on the exchange stream: after the owners' builds, after the launches that
        still read those slots; the launches enqueued later wait for it. No host synchronisation."""
        ses.external_write_begin(pictures, share["xs"].cuda_stream)
        with torch.cuda.stream(share["xs"]):
            share["moved"] = shard.all_gather_groups(groups, rank, world, dist, torch, share["pool"], share["slot_bytes"], share["cache"])
        ses.external_write_end(share["xs"].cuda_stream)

    def put_shared(base_ptr, device):
        """Owners upload + build, then every group of N slots is all-gathered in place."""
        ses.put_pictures(share["owned"], [base_ptr + f * frame_bytes for f in share["owned"]], W, device=device, bits=bits)
        groups = list(range(n_slots // world))
        exchange(groups, list(range(n_slots)))

    def step_resident():
        if share:
            put_shared(dev.data_ptr(), True)
        else:
            ses.put_pictures(frames_here, [dev.data_ptr() + f * frame_bytes for f in frames_here], W, device=True, bits=bits)
        ses.launch_jobs(jobs)

    # e2e: the look-ahead window is processed in decode order, chunk by chunk, through the queued form
    # of the C ABI: upload of chunk k+1 / pipeline of chunk k / download of chunk k-1 overlap. Every
    # frame is uploaded exactly once per step, right before the first chunk that needs it.
    sizes = [int(x) for x in args.e2e_chunks.split(",") if x]
    chunks, pos, have = [], 0, set()
    while pos < len(specs):
        n = sizes[min(len(chunks), len(sizes) - 1)]
        part = specs[pos:pos + n]
        pos += n
        need = []
        for _, cur, refs in part:
            for f in [cur] + list(refs.values()):
                if f not in have:
                    have.add(f)
                    need.append(f)
        cj, co = ses.make_jobs(part, deep=False, pin=args.pin_results)
        # the same chunk on the alternate slot set: picture numbers + ALT everywhere (distances are unchanged), same outputs
        alt = []
        for p0, cur, refs in part:
            p1 = abi.SvtMeParams.from_buffer_copy(bytes(p0))
            p1.picture_number = cur + ALT
            for (li, ri), n_ in refs.items():
                p1.ref_picture_number[li][ri] = n_ + ALT
            alt.append((p1, cur + ALT, {k: v + ALT for k, v in refs.items()}))
        cj_alt, co_alt = ses.make_jobs(alt, deep=False, pin=args.pin_results)
        chunks.append((need, [host_e2e.data_ptr() + f * e2e_frame_bytes for f in need], cj, co, cj_alt, co_alt))
    e2e_frames = len(share["owned"]) if share else len(have)
    chunks_sh = []
    if share:
        # pipelined picture-sharded run: GLOBAL chunks in decode order; every rank runs its share of each chunk and all
        # ranks exchange the same groups of ring slots before it (owners upload + build, one in-place all-gather per group)
        gsz = [world * int(x) for x in args.e2e_chunks_shared.split(",") if x]
        pos, have_g = 0, set()
        while pos < len(all_specs):
            n = gsz[min(len(chunks_sh), len(gsz) - 1)]
            idx = range(pos, min(pos + n, len(all_specs)))
            pos += n
            pics = [f for i in idx for f in [all_specs[i][1]] + list(all_specs[i][2].values())]
            groups = [g for g in shard.gather_groups(pics, world) if g not in have_g]
            have_g.update(groups)
            part = [all_specs[i] for i in idx if i % world == rank]
            cj, co = ses.make_jobs(part, deep=False, pin=args.pin_results) if part else (None, [])
            alt = []
            for p0, cur, refs in part:
                p1 = abi.SvtMeParams.from_buffer_copy(bytes(p0))
                p1.picture_number = cur + ALT
                for (li, ri), n_ in refs.items():
                    p1.ref_picture_number[li][ri] = n_ + ALT
                alt.append((p1, cur + ALT, {k: v + ALT for k, v in refs.items()}))
            cj_alt, co_alt = ses.make_jobs(alt, deep=False, pin=args.pin_results) if alt else (None, [])
            own = [g * world + rank for g in groups if g * world + rank < F]
            chunks_sh.append((groups, own, [host_e2e.data_ptr() + f * e2e_frame_bytes for f in own], cj, co, cj_alt, co_alt))
    sharing_info = None
    if share:
        # one-time check outside the timed region: a pyramid received over NCCL equals the one built locally
        put_shared(dev.data_ptr(), True)
        torch.cuda.synchronize()
        foreign = [p for p in frames_here if shard.pyramid_owner(p, world) != rank][:3]
        with me.MeSession(local_rank, W, H, capacity=1) as chk:
            for pnum in foreign:
                chk.put_pictures([pnum], [dev.data_ptr() + pnum * frame_bytes], W, device=True, bits=bits)
                for lvl in range(3):
                    if not np.array_equal(chk.get_plane(pnum, lvl), ses.get_plane(pnum, lvl)):
                        raise RuntimeError("rank %d: pyramid of picture %d (level %d) received over NCCL differs from a local build"
                                           % (rank, pnum, lvl))
                chk.release_picture(pnum)
        sharing_info = {"collective": "in-place ncclAllGather (torch.distributed.all_gather_into_tensor) over the N consecutive ring slots "
                                      "of a picture group, one per group, on an exchange stream ordered by events against the builds "
                                      "and the launches (svtme_external_write_begin / _end): no host synchronisation",
                        "groups_per_step": n_slots // world, "bytes_received_per_step": int(share["moved"]) if not chunks_sh else
                        int((world - 1) * share["slot_bytes"] * (n_slots // world)),
                        "pictures_built_here": len(share["owned"]), "verified_against_local_build": foreign,
                        "e2e_chunks": [len(c[4]) for c in chunks_sh]}

    # Results are collected (wait for the download, scatter into the callers' SvtMeOut arrays) by a second host thread, in
    # submission order, so that the thread that uploads / exchanges / submits never waits for a download: the reference's
    # ME threads and the downstream process that consumes MotionEstimationResults are different threads too
    # (me_process.c:304-313 posts to a FIFO). The step ends when the last batch has been collected.
    import queue
    collect_q = queue.Queue()
    collect_err = []
    # at most two steps' worth of batches in flight: the library recycles its batch objects (device + pinned result
    # buffers), so nothing is allocated inside a timed region once the warm-up has created that many
    in_flight = threading.BoundedSemaphore(2 * max(len(chunks_sh) if share else len(chunks), 1))

    def collector():
        while True:
            cj = collect_q.get()
            if cj is None:
                return
            try:
                if not collect_err:
                    ses.collect_jobs(cj)
            except Exception as e:  # noqa: BLE001
                collect_err.append(e)
            finally:
                in_flight.release()
                collect_q.task_done()

    collector_thread = threading.Thread(target=collector, daemon=True)
    collector_thread.start()

    def finish_step():
        collect_q.join()
        if collect_err:
            raise collect_err[0]

    def step_e2e_shared(drain=True, alt=False):
        n_groups = n_slots // world
        for groups, own, ptrs, cj, _, cj_alt, _ in chunks_sh:
            base = ALT if alt else 0
            if own:
                ses.put_pictures([base + f for f in own], ptrs, W, device=False, bits=e2e_bits)
            if groups:
                exchange([g + (n_groups if alt else 0) for g in groups], [base + p for g in groups for p in range(g * world, (g + 1) * world)])
            if cj is not None:
                in_flight.acquire()
                ses.submit_jobs(cj_alt if alt else cj)
                collect_q.put(cj_alt if alt else cj)
        if drain:
            finish_step()

    def step_e2e(drain=True, alt=False):
        if share:
            return step_e2e_shared(drain, alt)
        for c, (need, ptrs, cj, _, cj_alt, _) in enumerate(chunks):
            if alt:
                need, cj = [f + ALT for f in need], cj_alt
            if need:
                ses.put_pictures(need, ptrs, W, device=False, bits=e2e_bits)
            in_flight.acquire()
            ses.submit_jobs(cj)
            collect_q.put(cj)
        if drain:
            finish_step()

    def steps_e2e_streamed():
        """K steps as a continuous stream: the next step's uploads and launches are enqueued while the previous step's
        last batches still compute and download (no drain between steps, like pictures flowing through an encoder's
        look-ahead); the region ends when every result of every step has been collected."""
        for k in range(args.steps):
            step_e2e(drain=False, alt=bool(k & 1))
        finish_step()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    ses.set_search_timing(True)  # the search launches of each slot also get events of their own (roofline: the kernel's own duration)
    for _ in range(max(3, args.warmup)):
        step_resident()
    torch.cuda.synchronize()
    launches0 = me.launch_count()
    sampler = ClockSampler(local_rank)
    if rank == 0:  # only rank 0 reports; the other ranks keep their core
        sampler.start()
    ms_total = timed(step_resident, args.steps)
    launches = me.launch_count() - launches0
    # dominant kernel: the stage kernel of the last step (CUDA events inside the library, same stream)
    work_ops = ses.last_work_ops()
    kernel_ms = ses.last_kernel_ms()
    phases = ses.last_phase_ms()
    search_ms = ses.last_search_ms()
    phase_work = ses.last_phase_work()
    ses.set_search_timing(False)
    for _ in range(max(1, args.warmup)):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)
    steps_e2e_streamed()  # warm-up: creates the batch objects the deeper queue needs
    ms_e2e_streamed = timed(steps_e2e_streamed, 1)
    sampler.stop_flag = True
    if rank == 0:
        sampler.join(timeout=2)
    if args.timeline:
        # one more drained step, instrumented: host time (ms after the step started) at which each call returned, and the
        # device spans of each batch (pipeline start / end, download end; svtme_debug_last_span) measured from the same mark
        barrier()
        lib, spans, tl = ses.lib, [], []
        lib.svtme_debug_mark(ses.h)
        t0 = time.perf_counter()
        stamp = lambda what: tl.append((what, round((time.perf_counter() - t0) * 1e3, 3)))  # noqa: E731
        seq_c = share and chunks_sh or chunks
        for c, ch in enumerate(seq_c):
            if share:
                groups, own, ptrs, cj = ch[0], ch[1], ch[2], ch[3]
                if own:
                    ses.put_pictures(own, ptrs, W, device=False, bits=e2e_bits)
                    stamp("put%d(%d pics)" % (c, len(own)))
                if groups:
                    exchange(groups, [p for g in groups for p in range(g * world, (g + 1) * world)])
                    stamp("exchange%d(%d groups)" % (c, len(groups)))
            else:
                need, ptrs, cj = ch[0], ch[1], ch[2]
                if need:
                    ses.put_pictures(need, ptrs, W, device=False, bits=e2e_bits)
                    stamp("put%d(%d pics)" % (c, len(need)))
            if cj is not None:
                ses.submit_jobs(cj)
                stamp("submit%d(%d jobs)" % (c, len(cj)))
        for c, ch in enumerate(seq_c):
            cj = ch[3] if share else ch[2]
            if cj is None:
                continue
            ses.collect_jobs(cj)
            stamp("collect%d" % c)
            v = (C.c_float * 3)()
            lib.svtme_debug_last_span(ses.h, v)
            spans.append([round(float(x), 3) for x in v])
        if rank == 0:
            sys.stderr.write("TIMELINE host(ms): %s\nTIMELINE device spans per batch [pipeline start, pipeline end, download end] (ms): %s\n"
                             % (tl, spans))
        barrier()
    ms_step = ms_total / args.steps
    ms_e2e_best = min(ms_e2e_streamed, ms_e2e)  # two schedules of the same K steps and the same copies: the faster one is the headline
    sbs_step = (F - 1) * n_b64          # b64 of one whole stream
    n_streams = 1 if by_picture else world
    value = n_streams * sbs_step / (ms_step * 1e-3)
    e2e_val = n_streams * sbs_step / (ms_e2e / args.steps * 1e-3)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak, hbm_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback")
        vsad = float(ses.lib.svtme_measure_vabsdiff4_ginst_per_s())
        int_peak = vsad * 4 / 1e3  # T byte-SAD/s
        int_achieved = work_ops / (kernel_ms * 1e-3) / 1e12
        dom = max((k for k in phases if phase_work.get(k, 0) > 0), key=lambda k: phases[k])
        # the dominant KERNEL's own launch duration (events right before and after its launches, on the launching stream, in the
        # last timed step); the slot's time, which also holds the one-thread-per-b64 bookkeeping launches, is given beside it
        dom_ms = search_ms[dom] if search_ms.get(dom, 0) > 0 else phases[dom]
        dom_achieved = phase_work[dom] / (dom_ms * 1e-3) / 1e12
        slot_achieved = phase_work[dom] / (phases[dom] * 1e-3) / 1e12
        # DRAM traffic of the dominant launch comes from the committed `ncu --set full` capture of this
        # same command (profiles/): it cannot be measured outside the profiler
        traffic, traffic_src = None, None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            if tj.get("workload") == args.workload and tj.get("slot") == dom:
                traffic, traffic_src = tj["dram_bytes"], tj["source"]
        except Exception:
            pass
        # algorithmic HBM bytes of the stage kernel: every job reads its own pyramid and those of its
        # distinct references once (1 + 1/4 + 1/16 bytes per pixel) and writes its result tables
        alg_bytes = sum((1 + len(set(refs.values()))) * W * H * 1.3125 for _, _, refs in specs) + d2h_bytes
        cpu = cpu_reference_run(wl, specs, seq, cpu_pictures, threads, args.cpu_seconds)
        # bit-exactness of THIS run's GPU results against the reference outputs of the CPU sample: the arrays the
        # end-to-end steps left in the callers' SvtMeOut buffers, and the device-resident launch fetched now
        ses.fetch_jobs(jobs)
        e2e_outs = [o for c in chunks_sh for o in c[4]] if share else [o for c in chunks for o in c[3]]
        checks = [("e2e", e2e_outs), ("resident", outs)]
        if args.steps > 1:
            checks.append(("e2e streamed (alternate slot set)", [o for c in (chunks_sh if share else chunks) for o in c[6 if share else 5]]))
        n_arrays, bad = compare_outputs(cpu["outs"], checks)
        parity = {"pictures": len(cpu["outs"]), "arrays": n_arrays, "bit_exact": not bad, "against": "oracle/_ref (compiled "
                  "reference, %s kernels)" % cpu["simd"] if cpu["kind"] == "reference" else "oracle port",
                  "compared": "all SvtMeOut arrays of the sample's pictures, end-to-end and device-resident runs"}
        if bad:
            parity["mismatches"] = bad[:8]
        line = {"metric": "open-loop ME SBs/sec", "value": value, "unit": "SB/s", "n_gpus": world, "steps": args.steps,
                "warmup": max(3, args.warmup), "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "strong" if by_picture else "weak",
                "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": config,
                "frames_per_s": n_streams * F / (ms_step * 1e-3),
                # headline end-to-end number: the K steps as a continuous stream of look-ahead windows (host buffers in, host
                # result arrays out, every step's copies inside the timed region); the per-step-drained form is given beside it
                "e2e": {"value": n_streams * sbs_step / (ms_e2e_best / args.steps * 1e-3), "unit": "SB/s",
                        "h2d_bytes_per_step": e2e_frames * e2e_frame_bytes, "d2h_bytes_per_step": d2h_bytes,
                        "ms_per_step": ms_e2e_best / args.steps,
                        "frames_per_s": n_streams * F / (ms_e2e_best / args.steps * 1e-3),
                        "schedule": "streamed" if ms_e2e_streamed <= ms_e2e else "drained (faster than the streamed schedule in this run)",
                        "streamed_ms_per_step": ms_e2e_streamed / args.steps,
                        "mode": "streamed: the %d timed steps are enqueued back to back (consecutive steps use alternating sets of "
                                "ring slots, like new pictures of a stream), results are collected in order by a second host thread, "
                                "the region ends when the last result array is filled" % args.steps,
                        "api": ("svtme_put_pictures_10bit (pinned host 16-bit luma)" if e2e_bits == 10 else
                                ("svtme_put_pictures (pinned host 8-bit MSB plane of the 10-bit source, the buffer the reference's ME reads)"
                                 if bits == 10 else "svtme_put_pictures (pinned host luma)")) +
                               " + svtme_submit_pictures / svtme_collect_results (results in host SvtMeOut arrays%s), batches of %s pictures"
                               % (", the three big tables page-locked and filled by the download itself" if args.pin_results else "",
                                  args.e2e_chunks_shared + " per rank" if share else args.e2e_chunks),
                        "drained": {"value": e2e_val, "unit": "SB/s", "ms_per_step": ms_e2e / args.steps,
                                    "note": "every step drained (all results collected) before the next one starts"}},
                "gpu_launches": int(launches),
                **({"pyramid_sharing": sharing_info} if sharing_info else {}),
                # the dominant kernel of the pipeline, timed by itself; `slot` = the launch slot it runs in (the kernel + the scalar
                # plan/post launch(es) after it); the whole pipeline is given next to it
                "roofline": {"bound": "int_alu", "kernel": "%s, launch slot '%s' (the kernel %.0f%%, the slot %.0f%% of the pipeline time)"
                                                           % (SLOT_KERNELS.get(dom, "stage pipeline"), dom, 100 * dom_ms / max(kernel_ms, 1e-9),
                                                              100 * phases[dom] / max(kernel_ms, 1e-9)),
                             "achieved": dom_achieved, "peak": int_peak, "unit": "Tbyte-SAD/s",
                             "frac": dom_achieved / int_peak if int_peak else None, "traffic": traffic,
                             "traffic_source": traffic_src,
                             "peak_source": "VABSDIFF4.U8.ACC issue rate measured live in this run (%.0f Ginst/s x 4 bytes)" % vsad,
                             "work_byte_ops_per_launch": phase_work[dom], "launch_ms": dom_ms,
                             "slot": {"ms": phases[dom], "achieved": slot_achieved, "frac": slot_achieved / int_peak if int_peak else None,
                                      "note": "kernel + the bookkeeping launches timed in the same slot"},
                             "search_ms": {k: round(v, 3) for k, v in search_ms.items() if v > 0},
                             "pipeline": {"achieved": int_achieved, "frac": int_achieved / int_peak if int_peak else None,
                                          "work_byte_ops": work_ops, "ms": kernel_ms, "share_of_step": kernel_ms / ms_step},
                             "phase_ms": {k: round(v, 3) for k, v in phases.items()},
                             "phase_frac": {k: round(phase_work[k] / (phases[k] * 1e-3) / 1e12 / int_peak, 4)
                                            for k in phases if phases[k] > 0 and phase_work[k] > 0}},
                "roofline_hbm": {"bound": "hbm", "kernel": "stage pipeline", "achieved": alg_bytes / (kernel_ms * 1e-3) / 1e9,
                                 "peak": hbm_peak, "unit": "GB/s", "frac": alg_bytes / (kernel_ms * 1e-3) / 1e9 / hbm_peak,
                                 "traffic": None, "peak_source": hbm_src},
                "cpu_baseline": {"value": cpu["sbs_per_s"], "unit": "SB/s", "cores": threads, "kind": cpu["kind"],
                                 "frames_per_s": cpu["fps"], "parallel_efficiency": cpu["parallel_efficiency"],
                                 "value_1thread": cpu["sbs_per_s_1thread"], "sample": cpu_sample_text(cpu)},
                "parity": parity,
                "clocks": sampler.summary()}
        print(json.dumps(line))
        if bad:
            sys.stderr.write("bench: GPU results differ from the reference: %s\n" % ", ".join(bad[:8]))
            exit_code = 1
    collect_q.put(None)
    ses.close()
    if world > 1:
        dist.destroy_process_group()
    return exit_code


if __name__ == "__main__":
    sys.exit(main())
