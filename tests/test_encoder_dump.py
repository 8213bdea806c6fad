"""Encoder-level parity: fixtures dumped from the REAL reference encoder (SvtAv1EncApp built from
the reference with two hook calls around svt_aom_motion_estimation_b64, oracle/encoder_dump/, generator
tests/golden/make_encoder_dump.py) pin (a) the host-side parameter derivation against what the encoder's own
MeContext held for every ME picture (enc_mode_config.c:684-823 as driven by pd_process.c / me_process.c:219-262),
(b) the prediction-structure facts the path depends on, (c) the oracle and (d, -m gpu) the CUDA path against the
MeSbResults and per-b64 pcs arrays the encoder really produced, and (e, -m gpu) the bitstream: the encoder fed with
the GPU's ME results writes the same .ivf, byte for byte."""
import ctypes as C
import os
import subprocess
import sys
import zlib

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import _encdump as E
import _util as U
import make_encoder_dump as M

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = [c[0] for c in M.CASES]
SLOT_KEYS = {"hme_sc": 2, "hme_sad": 1, "do_ref": 1}  # per (list, ref) state: slots the picture does not use are stale in the encoder


def load_case(name):
    z = np.load(os.path.join(GOLD, "encdump_%s.npz" % name))
    w, h, n, preset, deep = (int(v) for v in z["meta"])
    pics = []
    for pic in z["pictures"].tolist():
        hdr = E.DumpHeader.from_buffer_copy(z["hdr_%d" % pic].tobytes())
        arrays = {k[:-len("_%d" % pic)]: z[k] for k in z.files if k.endswith("_%d" % pic) and not k.startswith("hdr_")}
        pics.append((pic, hdr, arrays))
    seq = U.synth_sequence(w, h, n, seed=1)
    for pic, crc in zip(z["luma_pictures"].tolist(), z["luma_crc32"].tolist()):
        assert zlib.crc32(seq[pic].tobytes()) == crc, "synthetic frame %d is not what the encoder's ME read" % pic
    return w, h, seq, pics


def refs_of(hdr):
    p = hdr.params
    return {(li, ri): int(hdr.ref_picture_number[li][ri]) for li in range(p.num_of_list_to_search)
            for ri in range(p.num_of_ref_pic_to_search[li])}


def assert_outputs_equal(hdr, want, got, what):
    used = np.zeros((2, 4), dtype=bool)
    for (li, ri) in refs_of(hdr):
        used[li, ri] = True
    n_b64 = hdr.b64_cols * hdr.b64_rows
    for k, w in want.items():
        g = got[k]
        if k in SLOT_KEYS:
            w = w.reshape(n_b64, 2, 4, -1)[:, used]
            g = g.reshape(n_b64, 2, 4, -1)[:, used]
        if k == "me_candidate_array":  # entries beyond total_me_candidate_index are leftovers of earlier pictures in the encoder
            tot = np.asarray(want["total_me_candidate_index"]).reshape(n_b64, hdr.n_pu)
            live = np.arange(hdr.max_cand)[None, None, :] < tot[:, :, None]
            w = np.asarray(w).reshape(n_b64, hdr.n_pu, hdr.max_cand)[live]
            g = np.asarray(g).reshape(n_b64, hdr.n_pu, hdr.max_cand)[live]
        if k == "me_mv_array":
            # [pu][max_refs]: the encoder writes a slot only where a candidate of the PU may use it (construct_me_candidate_array,
            # motion_estimation.c:2658-2695: a reference pruned for the PU is skipped) and recycles the buffers between
            # pictures, so the other slots hold whatever an earlier picture left; only referenced slots are results
            cand = np.asarray(want["me_candidate_array"]).reshape(n_b64, hdr.n_pu, hdr.max_cand)
            tot = np.asarray(want["total_me_candidate_index"]).reshape(n_b64, hdr.n_pu)
            live = np.arange(hdr.max_cand)[None, None, :] < tot[:, :, None]
            mask = np.zeros((n_b64, hdr.n_pu, hdr.max_refs), dtype=bool)
            d, i0, i1, l0, l1 = cand & 3, (cand >> 2) & 3, (cand >> 4) & 3, (cand >> 6) & 1, (cand >> 7) & 1
            bi, pi = np.meshgrid(np.arange(n_b64), np.arange(hdr.n_pu), indexing="ij")
            for c in range(hdr.max_cand):
                for use, lst, idx in (((d[:, :, c] == 0) | (d[:, :, c] == 2), l0[:, :, c], i0[:, :, c]),
                                      ((d[:, :, c] == 1) | (d[:, :, c] == 2), l1[:, :, c], i1[:, :, c])):
                    sel = live[:, :, c] & use
                    slot = np.where(lst == 1, hdr.max_l0, 0) + idx
                    ok = sel & (slot < hdr.max_refs)
                    mask[bi[ok], pi[ok], slot[ok]] = True
            assert mask.any()
            w = np.asarray(w).reshape(n_b64, hdr.n_pu, hdr.max_refs)[mask]
            g = np.asarray(g).reshape(n_b64, hdr.n_pu, hdr.max_refs)[mask]
        assert np.array_equal(np.asarray(w).reshape(-1), np.asarray(g).reshape(-1)), (what, int(hdr.picture_number), k)


@pytest.mark.parametrize("name", CASES)
def test_derived_parameters_equal_the_encoders_me_context(name):
    """(a) svtme_derive_params(SvtMeDeriveInput as the encoder's scs/pcs had it) + the per-picture fields ==
    the SvtMeParams rebuilt from the MeContext the encoder handed to svt_aom_motion_estimation_b64."""
    lib = U.product_lib()
    w, h, seq, pics = load_case(name)
    assert len(pics) >= 5
    for pic, hdr, _ in pics:
        q = U.abi.SvtMeParams()
        assert lib.svtme_derive_params(C.byref(hdr.din), C.byref(q)) == 0
        U.finish_params(q, hdr.din, pic, refs_of(hdr))
        assert E.struct_fields_equal(hdr.params, q) == [], (name, pic)
        assert lib.svtme_exported_pu_count(C.byref(q)) == hdr.n_pu


def test_prediction_structure_facts_the_path_relies_on():
    """(b) what pd_process.c really posts as PAME tasks (the job planner gop.py is only a workload generator): every
    non-intra picture is ME'd once; references are source pictures that precede the picture in decode order; base
    layer pictures search list 0 only with two lists declared; list sizes stay within the allocation."""
    for name in CASES:
        w, h, seq, pics = load_case(name)
        order = {pic: hdr.decode_order for pic, hdr, _ in pics}
        order[0] = 0
        assert sorted(order) == list(range(len(order)))
        for pic, hdr, _ in pics:
            p = hdr.params
            for (li, ri), r in refs_of(hdr).items():
                assert order[r] < order[pic], (name, pic, r)
            assert p.num_of_ref_pic_to_search[0] <= p.max_l0 and p.num_of_ref_pic_to_search[0] + p.num_of_ref_pic_to_search[1] <= p.max_refs
            if p.temporal_layer_index == 0:
                assert all(r < pic for r in refs_of(hdr).values())


@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_the_encoders_me_results(name):
    """(c) the CPU checkers, run on the synthetic frames with the dumped parameters, give the arrays the encoder dumped."""
    w, h, seq, pics = load_case(name)
    n_b64 = ((w + 63) // 64) * ((h + 63) // 64)
    checkers = ["orc"] + (["ref"] if U.ref_lib() is not None else [])
    for which in checkers:
        if which == "ref":
            U.ref_lib().ref_set_simd(1)
        pyrs = U.CpuPyramidSet(which).build({n: seq[n] for n in range(len(seq))}, 4)
        specs = [(hdr.params, pic, refs_of(hdr)) for pic, hdr, _ in pics]
        jobs, outs = U.cpu_mt_jobs(which, specs, pyrs, n_b64, deep=True)
        U.cpu_me_pictures_mt(which, jobs, 8)
        for (pic, hdr, want), got in zip(pics, outs):
            assert_outputs_equal(hdr, want, got, which)
        pyrs.free()


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_gpu_reproduces_the_encoders_me_results(name):
    """(d) the CUDA path through the C ABI: parameters from svtme_derive_params, pictures from the synthetic clip."""
    import importlib
    me = importlib.import_module(U.PKG + ".me")
    w, h, seq, pics = load_case(name)
    with me.MeSession(0, w, h, capacity=len(seq)) as ses:
        if name.endswith("10bit"):  # the 16-bit samples the encoder was fed: the boundary keeps the 8 MSBs, as the encoder's ME did
            host = np.ascontiguousarray(M.samples16(seq))
            ses.put_pictures(list(range(len(seq))), [host[i].ctypes.data for i in range(len(seq))], w, bits=10)
        else:
            host = np.ascontiguousarray(seq)
            ses.put_pictures(list(range(len(seq))), [host[i].ctypes.data for i in range(len(seq))], w)
        specs = []
        for pic, hdr, _ in pics:
            q = ses.derive_params(hdr.din)
            specs.append((U.finish_params(q, hdr.din, pic, refs_of(hdr)), pic, refs_of(hdr)))
        jobs, outs = ses.make_jobs(specs, deep=True)
        ses.run_jobs(jobs)
    for (pic, hdr, want), got in zip(pics, outs):
        assert_outputs_equal(hdr, want, got, "gpu")


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(M.APP), reason="oracle/_ref/SvtAv1EncApp_dump not built (needs /root/reference once)")
@pytest.mark.parametrize("name,w,h,n,preset,extra", [("360p_p8", 640, 360, 16, 8, []), ("360p_p4", 640, 360, 10, 4, []),
                                                     ("360p_p9_ld", 640, 360, 9, 9, ["--pred-struct", "1"])])
def test_bitstream_is_identical_with_gpu_me_results(name, w, h, n, preset, extra, tmp_path):
    """(e) north_star "identical bitstream": the reference encoder is run (1) untouched, (2) with every b64's ME
    results REPLACED by the GPU's (SVTME_REPLAY_DIR), (3) with one GPU motion vector falsified. (1) and (2) must be
    the same file; (3) must not (the replacement really reaches mode decision)."""
    import importlib
    me = importlib.import_module(U.PKG + ".me")
    seq = U.synth_sequence(w, h, n, seed=1)
    yuv = str(tmp_path / "in.yuv")
    M.write_yuv(yuv, seq)
    ddir, gdir, xdir = tmp_path / "dump", tmp_path / "gpu", tmp_path / "bad"
    for d in (ddir, gdir, xdir):
        d.mkdir()
    M.encode(yuv, w, h, n, preset, extra, str(tmp_path / "ref.ivf"), {"SVTME_DUMP_DIR": str(ddir)})
    pics = sorted(int(f[3:9]) for f in os.listdir(ddir) if f.startswith("me_"))
    with me.MeSession(0, w, h, capacity=n) as ses:
        host = np.ascontiguousarray(seq)
        ses.put_pictures(list(range(n)), [host[i].ctypes.data for i in range(n)], w)
        hdrs, specs = [], []
        for pic in pics:
            hdr, _ = E.read_dump(str(ddir / ("me_%06d.bin" % pic)))
            q = ses.derive_params(hdr.din)
            hdrs.append(hdr)
            specs.append((U.finish_params(q, hdr.din, pic, refs_of(hdr)), pic, refs_of(hdr)))
        jobs, outs = ses.make_jobs(specs, deep=False)
        ses.run_jobs(jobs)
    for pic, hdr, got in zip(pics, hdrs, outs):
        E.write_dump(str(gdir / ("me_%06d.bin" % pic)), hdr, got)
        bad = {k: v.copy() for k, v in got.items()}
        if pic == pics[len(pics) // 2]:
            bad["me_mv_array"].reshape(-1)[::7] ^= 0x00040008  # falsify motion vectors of one picture
        E.write_dump(str(xdir / ("me_%06d.bin" % pic)), hdr, bad)
    M.encode(yuv, w, h, n, preset, extra, str(tmp_path / "gpu.ivf"), {"SVTME_REPLAY_DIR": str(gdir)})
    M.encode(yuv, w, h, n, preset, extra, str(tmp_path / "bad.ivf"), {"SVTME_REPLAY_DIR": str(xdir)})
    ref, gpu, bad = (open(str(tmp_path / f), "rb").read() for f in ("ref.ivf", "gpu.ivf", "bad.ivf"))
    assert len(ref) > 1000
    assert ref == gpu, "bitstream written with the GPU's ME results differs from the reference encoder's"
    assert ref != bad, "falsified ME results did not change the bitstream: the replay hook is not live"


# ---------------------------------------------------------------- next to ME: picture statistics and the DG detector
AUX_CASES = [c for c in CASES + [a[0] for a in M.AUX_ONLY] if os.path.exists(os.path.join(GOLD, "encdump_aux_%s.npz" % c))]


def load_aux(name):
    z = np.load(os.path.join(GOLD, "encdump_aux_%s.npz" % name))
    w, h, n, preset = (int(v) for v in z["meta"])
    return z, w, h, U.synth_sequence(w, h, n, seed=1)


def stats_params_of(z, pic):
    v = [int(x) for x in z["params_%d" % pic]]  # calc_hist, calculate_variance, prec_sub, all_sizes, scd, regions w / h
    return U.stats_params(calc_hist=v[0], calculate_variance=v[1], prec_sub=v[2], all_sizes=v[3], scd=v[4], regions=(v[5], v[6]))


def assert_stats_equal(z, pic, p, got, what):
    keys = ["pic_avg_variance", "avg_luma"] + (["variance"] if p.calculate_variance else []) + \
        (["picture_histogram", "average_intensity_per_region"] if p.calc_hist else [])
    for k in keys:
        assert np.array_equal(z["%s_%d" % (k, pic)], got[k]), (what, pic, k)


def test_aux_fixtures_exist():
    assert len(AUX_CASES) >= 3 and "dg_360p_p8" in AUX_CASES, "tests/golden/encdump_aux_*.npz missing (tests/golden/make_encoder_dump.py writes them)"


@pytest.mark.parametrize("name", AUX_CASES)
def test_oracle_reproduces_the_encoders_picture_statistics_and_dg_metrics(name):
    """What the real encoder's picture-analysis kernel left in pcs->variance / picture_histogram / avg_luma ... and what its
    dynamic-GOP detector summed up, against the CPU checkers run on the synthetic frames."""
    z, w, h, seq = load_aux(name)
    checkers = ["orc"] + (["ref"] if U.ref_lib() is not None else [])
    for which in checkers:
        pyr = {}
        for pic in z["stats_pictures"].tolist():
            pyr[pic] = U.CpuPyramid(which, seq[pic], pic)
            p = stats_params_of(z, pic)
            assert_stats_equal(z, pic, p, U.cpu_picture_statistics(which, pyr[pic], p), which)
        for src, ref, dist, cplx, active, inv, res in z["dg"].tolist():
            for pic in (src, ref):
                if pic not in pyr:
                    pyr[pic] = U.CpuPyramid(which, seq[pic], pic)
            assert U.cpu_dg_detector_hme(which, pyr[src], pyr[ref], res) == (dist, cplx, active, inv), (which, src, ref)


@pytest.mark.gpu
@pytest.mark.parametrize("name", AUX_CASES)
def test_gpu_reproduces_the_encoders_picture_statistics_and_dg_metrics(name):
    import importlib
    me = importlib.import_module(U.PKG + ".me")
    z, w, h, seq = load_aux(name)
    n_b64 = ((w + 63) // 64) * ((h + 63) // 64)
    with me.MeSession(0, w, h, capacity=len(seq)) as ses:
        host = np.ascontiguousarray(seq)
        ses.put_pictures(list(range(len(seq))), [host[i].ctypes.data for i in range(len(seq))], w)
        for pic in z["stats_pictures"].tolist():
            p = stats_params_of(z, pic)
            out, arrays = U.abi.alloc_stats_out(n_b64, p)
            assert ses.lib.svtme_picture_statistics(ses.h, pic, C.byref(p), C.byref(out)) == 0, ses.lib.svtme_last_error()
            assert_stats_equal(z, pic, p, arrays, "gpu")
        for src, ref, dist, cplx, active, inv, res in z["dg"].tolist():
            m = U.abi.SvtMeDgMetrics()
            assert ses.lib.svtme_dg_detector_hme(ses.h, src, ref, res, C.byref(m)) == 0, ses.lib.svtme_last_error()
            assert (int(m.tot_dist), int(m.tot_cplx), int(m.tot_active), int(m.sum_in_vectors)) == (dist, cplx, active, inv), (src, ref)


# ---------------------------------------------------------------- temporal-filter ME calls of the real encoder
TF_CASES = [c[0] for c in M.TF_CASES if os.path.exists(os.path.join(GOLD, "encdump_%s.npz" % c[0]))]


def load_tf(name):
    z = np.load(os.path.join(GOLD, "encdump_%s.npz" % name))
    w, h, n, preset = (int(v) for v in z["meta"])
    pairs = []
    for c, f in z["pairs"].tolist():
        hdr = E.TfDumpHeader.from_buffer_copy(z["hdr_%d_%d" % (c, f)].tobytes())
        pairs.append((c, f, hdr, {k: z["%s_%d_%d" % (k, c, f)] for k in ("p_sb_best_sad", "p_sb_best_mv", "hme_sc", "hme_sad")}))
    return w, h, U.synth_sequence(w, h, n, seed=1), pairs


def assert_tf_outputs(hdr, want, got, what):
    """Slot (list 0, ref 0) of the deep arrays; a block that returned early after HME keeps stale tables in the encoder."""
    n = hdr.b64_cols * hdr.b64_rows
    early = want["hme_sad"] < hdr.params.tf_me_exit_th
    g_sad = np.asarray(got["hme_sad"]).reshape(n, 2, 4)[:, 0, 0]
    g_sc = np.asarray(got["hme_sc"]).reshape(n, 2, 4, 2)[:, 0, 0]
    assert np.array_equal(want["hme_sad"], g_sad), (what, "hme_sad")
    assert np.array_equal(want["hme_sc"], g_sc), (what, "hme_sc")
    for k in ("p_sb_best_sad", "p_sb_best_mv"):
        g = np.asarray(got[k]).reshape(n, 2, 4, 85)[:, 0, 0]
        assert np.array_equal(want[k][~early], g[~early]), (what, k)


def test_tf_fixtures_exist():
    assert len(TF_CASES) == 2


@pytest.mark.parametrize("name", TF_CASES)
def test_tf_parameters_equal_the_encoders_me_context(name):
    """svtme_derive_params_tf(what the encoder's pcs / scs held) + the per-call fields == the SvtMeParams rebuilt from the
    MeContext the temporal filter handed to svt_aom_motion_estimation_b64."""
    lib = U.product_lib()
    w, h, seq, pairs = load_tf(name)
    assert len(pairs) >= 8
    for c, f, hdr, _ in pairs:
        q = U.abi.SvtMeParams()
        assert lib.svtme_derive_params_tf(C.byref(hdr.tin), C.byref(q)) == 0
        q.picture_number, q.ref_picture_number[0][0] = c, f
        assert E.struct_fields_equal(hdr.params, q, skip=TF_NOT_IN_ME_CONTEXT) == [], (name, c, f)


# fields of SvtMeParams that live in the pcs / scs, not in MeContext, and only steer the candidate export that ME_MCTF skips
# (the dump hook of the temporal filter rebuilds the struct from the MeContext alone and leaves them zero)
TF_NOT_IN_ME_CONTEXT = ("enable_me_8x8", "enable_me_16x16", "max_number_of_pus_per_sb", "input_resolution_le_480p", "max_cand",
                        "max_refs", "max_l0")


def tf_params(hdr, c, f):
    q = U.abi.SvtMeParams()
    assert U.product_lib().svtme_derive_params_tf(C.byref(hdr.tin), C.byref(q)) == 0
    q.picture_number, q.ref_picture_number[0][0] = c, f
    return q


@pytest.mark.parametrize("name", TF_CASES)
def test_oracle_reproduces_the_encoders_tf_me_results(name):
    w, h, seq, pairs = load_tf(name)
    for which in ["orc"] + (["ref"] if U.ref_lib() is not None else []):
        pyr = {}
        for c, f, hdr, want in pairs:
            for pic in (c, f):
                if pic not in pyr:
                    pyr[pic] = U.CpuPyramid(which, seq[pic], pic)
            got = U.cpu_me_picture(which, tf_params(hdr, c, f), pyr[c], {(0, 0): pyr[f]})
            assert_tf_outputs(hdr, want, got, (which, c, f))


@pytest.mark.gpu
@pytest.mark.parametrize("name", TF_CASES)
def test_gpu_reproduces_the_encoders_tf_me_results(name):
    import importlib
    me = importlib.import_module(U.PKG + ".me")
    w, h, seq, pairs = load_tf(name)
    with me.MeSession(0, w, h, capacity=len(seq)) as ses:
        host = np.ascontiguousarray(seq)
        ses.put_pictures(list(range(len(seq))), [host[i].ctypes.data for i in range(len(seq))], w)
        specs = []
        for c, f, hdr, _ in pairs:
            q = U.abi.SvtMeParams()
            assert ses.lib.svtme_derive_params_tf(C.byref(hdr.tin), C.byref(q)) == 0
            q.picture_number, q.ref_picture_number[0][0] = c, f
            specs.append((q, c, {(0, 0): f}))
        jobs, outs = ses.make_jobs(specs, deep=True)
        ses.run_jobs(jobs)
    for (c, f, hdr, want), got in zip(pairs, outs):
        assert_tf_outputs(hdr, want, got, ("gpu", c, f))
