"""GPU parity at the BASELINE.json picture sizes (-m gpu): whole ME pictures of the bench workloads
(1080p preset 8, 4K 10-bit preset 4, 8K preset 6) and long one-line pictures, through the C ABI, against
the CPU checker (the compiled reference with its AVX2 kernels when oracle/_ref is present, else the
oracle restatement), bit for bit, deep state included. These are the first pictures with more than 32
b64 in a line, i.e. with more than one segment per line in strip_kernel (svtme_stage.cu, STRIP_SEG).
Nothing here reads /root/reference."""
import ctypes as C
import importlib

import numpy as np
import pytest

import _util as U

pytestmark = pytest.mark.gpu
me = importlib.import_module(U.PKG + ".me")
gop = importlib.import_module(U.PKG + ".gop")
THREADS = 8


def product_params(d):
    p = U.abi.SvtMeParams()
    assert U.product_lib().svtme_derive_params(C.byref(d), C.byref(p)) == 0
    return p


def workload_specs(W, H, preset, frames, levels, n_l0, n_l1, picks):
    specs = []
    jobs = gop.ra_jobs(frames, levels, n_l0, n_l1)
    for i in picks:
        j = jobs[i]
        d = U.derive_input(preset, W, H, hierarchical_levels=levels, temporal_layer_index=j.temporal_layer, is_ref=int(j.is_ref),
                           ref_list0_count_try=j.n_l0, ref_list1_count_try=j.n_l1, ref_count_used_list0=n_l0,
                           ref_count_used_list1=n_l1)
        specs.append((U.finish_params(product_params(d), d, j.picture_number, j.refs), j.picture_number, j.refs))
    return specs


def check_against_cpu(seq, specs, W, H, bits=8, deep=True, one_batch=True):
    which = U.best_cpu_checker()
    need = sorted({n for _, cur, refs in specs for n in [cur] + list(refs.values())})
    n_b64 = ((W + 63) // 64) * ((H + 63) // 64)
    pyrs = U.CpuPyramidSet(which).build({n: seq[n] for n in need}, THREADS)
    cj, want = U.cpu_mt_jobs(which, specs, pyrs, n_b64, deep=deep)
    U.cpu_me_pictures_mt(which, cj, THREADS)
    with me.MeSession(0, W, H, capacity=len(need)) as ses:
        if bits == 10:  # 16-bit samples: the 8-bit pattern << 2 plus two random LSBs (svtme_put_pictures_10bit)
            rng = np.random.default_rng(3)
            src = [np.ascontiguousarray((seq[n].astype(np.uint16) << 2) | rng.integers(0, 4, size=seq[n].shape, dtype=np.uint16))
                   for n in need]
        else:
            src = [np.ascontiguousarray(seq[n]) for n in need]
        ses.put_pictures(need, [s.ctypes.data for s in src], W, bits=bits)
        if one_batch:
            gj, got = ses.make_jobs(specs, deep=deep)
            ses.run_jobs(gj)
        else:
            got = [ses.run_picture(p, cur, refs, deep=deep) for p, cur, refs in specs]
        for lvl in range(3):  # the pyramid of the first picture too, pads included
            plane = ses.get_plane(need[0], lvl)
            cpu = np.ctypeslib.as_array(C.cast(getattr(pyrs[need[0]], ("full", "quarter", "sixteenth")[lvl]).buffer_y, C.POINTER(C.c_uint8)),
                                        shape=(plane.shape[0], plane.shape[1]))
            assert np.array_equal(cpu, plane), ("pyramid level", lvl)
    for (p, cur, refs), w, g in zip(specs, want, got):
        for k in w:
            if not np.array_equal(w[k], g[k]):
                bad = np.argwhere(w[k].reshape(n_b64, -1) != g[k].reshape(n_b64, -1))
                raise AssertionError("picture %d: %s differs in %d entries, first at b64 %d (col %d, row %d) entry %d"
                                     % (cur, k, len(bad), bad[0][0], bad[0][0] % ((W + 63) // 64), bad[0][0] // ((W + 63) // 64), bad[0][1]))
    pyrs.free()


def test_1080p_preset8_pictures_of_the_bench_workload():
    """BASELINE configs[1]: 1920x1080, preset 8, refs 3+2, five hierarchical levels; 30x17 b64, last row 56 high."""
    W, H = 1920, 1080
    seq = U.synth_sequence(W, H, 33, seed=1)
    specs = workload_specs(W, H, 8, 33, 5, 3, 2, [0, 1, 2, 5, 9, 17, 31])  # base layer, every other layer
    check_against_cpu(seq, specs, W, H)


def test_4k_preset4_10bit_pictures_of_the_bench_workload():
    """BASELINE configs[2]: 3840x2160 10-bit sources, preset 4, refs 4+3, HME level 2 on, full-pel areas up to 64x32;
    60x34 b64: two segments per b64 row and per b64 column in strip mode."""
    W, H = 3840, 2160
    seq = U.synth_sequence(W, H, 33, seed=1)
    specs = workload_specs(W, H, 4, 33, 5, 4, 3, [0, 1, 3, 10])
    check_against_cpu(seq, specs, W, H, bits=10)


def test_8k_preset6_pictures_of_the_bench_workload():
    """BASELINE configs[3]: 7680x4320, preset 6, mini-GOP 16, refs 3+2; 120x68 b64: four / three segments per line."""
    W, H = 7680, 4320
    seq = U.synth_sequence(W, H, 17, seed=1)
    specs = workload_specs(W, H, 6, 17, 4, 3, 2, [0, 2])
    check_against_cpu(seq, specs, W, H, deep=False)


@pytest.mark.parametrize("size", [(2200, 72), (72, 2200), (4160, 136)], ids=lambda s: "%dx%d" % s)
@pytest.mark.parametrize("preset", [8, 4, 1, 11])  # 11: pre-HME skips even search rows, but not for the narrow last column
def test_long_line_pictures_strip_segments(size, preset):
    """One or two lines of 35 / 65 b64 (ragged last segment: 35 = 32 + 3, 65 = 2 x 32 + 1; last column 24 wide, last row 8
    high): pre-HME and HME level 0 in strip mode with bx0 > 0 / by0 > 0, several references, motion into the pads."""
    W, H = size
    seq = U.synth_sequence(W, H, 7, seed=preset + W, motion=(7, -5) if W > H else (-3, 9), noise=3.0)
    refs, cur = {(0, 0): 2, (0, 1): 1, (0, 2): 0, (1, 0): 4, (1, 1): 5}, 3
    specs = []
    for tl, c, r in [(2, cur, refs), (0, 6, {(0, 0): 5, (0, 1): 3, (1, 0): 5, (1, 1): 3}), (3, 1, {(0, 0): 0, (1, 0): 2})]:
        l0 = 1 + max(ri for (li, ri) in r if li == 0)
        l1 = 1 + max(ri for (li, ri) in r if li == 1)
        d = U.derive_input(preset, W, H, temporal_layer_index=tl, ref_list0_count_try=l0, ref_list1_count_try=l1,
                           ref_count_used_list0=3, ref_count_used_list1=2)
        specs.append((U.finish_params(product_params(d), d, c, r), c, r))
    check_against_cpu(seq, specs, W, H)
    check_against_cpu(seq, specs[:1], W, H, one_batch=False)


def test_many_small_batches_of_different_pictures_in_flight():
    """Queued form under stress: 18 batches of one or two pictures each — presets 4 and 8 alternating (different kernels,
    reference counts and job descriptors) — in flight on the library's two compute streams, three rounds. CTAs of different
    batches then interleave on every SM; each must work from its own job descriptor and state. (The one bug of this kind so
    far — a kernel that read its descriptor from shared memory before the staging copy had landed — only showed in the queued
    4K run of bench.py, whose own parity check and tools/repro_e2e.py cover that size.)"""
    W, H = 640, 360
    seq = U.synth_sequence(W, H, 9, seed=5, noise=3.0)
    groups = []
    for preset, n_l0, n_l1 in ((4, 2, 2), (8, 2, 1)):
        specs = workload_specs(W, H, preset, 9, 3, n_l0, n_l1, list(range(8)))
        groups.append(specs)
    with me.MeSession(0, W, H, capacity=9) as ses:
        ses.put_pictures(list(range(9)), [np.ascontiguousarray(f).ctypes.data for f in seq], W)
        want = []
        for specs in groups:
            jobs, outs = ses.make_jobs(specs, deep=True)
            ses.run_jobs(jobs)
            want.append(outs)
        for rnd in range(3):
            queued = []
            pos = [0, 0]
            k = 0
            while pos[0] < 8 or pos[1] < 8:
                g = k % 2 if pos[k % 2] < 8 else 1 - k % 2
                n = 1 + (k + rnd) % 2
                part = groups[g][pos[g]:pos[g] + n]
                jobs, outs = ses.make_jobs(part, deep=True)
                ses.submit_jobs(jobs)
                queued.append((jobs, outs, g, pos[g]))
                pos[g] += len(part)
                k += 1
            for jobs, outs, g, first in queued:
                ses.collect_jobs(jobs)
                for i, o in enumerate(outs):
                    for name in o:
                        assert np.array_equal(o[name], want[g][first + i][name]), (rnd, g, first + i, name)
