"""Random stage configurations for differential tests: preset, picture size (complete and incomplete b64 in both directions),
temporal layer, reference counts and placement, encoder flags, content (translating texture with noise, static clip, pure
noise). Used by tests/test_fuzz.py (a small seeded sample) and tools/fuzz_*.py (as many as wanted). Test infrastructure only."""
import numpy as np

import _util as U


def random_config(rng):
    """-> (SvtMeDeriveInput, sequence, cur, refs, description) or None when the reference rejects the combination."""
    preset = int(rng.integers(-1, 14))
    W = int(rng.integers(2, 9)) * 64 + int(rng.choice([0, 8, 24, 40, 56]))
    H = int(rng.integers(2, 6)) * 64 + int(rng.choice([0, 8, 16, 32, 48]))
    tl = int(rng.integers(0, 4))
    levels = int(rng.integers(max(tl, 1), 6))
    l0 = int(rng.integers(1, 5))
    l1 = int(rng.integers(0, 4)) if tl > 0 or rng.random() < 0.5 else 0
    kw = dict(temporal_layer_index=tl, hierarchical_levels=levels, ref_list0_count_try=l0, ref_list1_count_try=l1, slice_type_b=int(l1 > 0),
              ref_count_used_list0=max(l0, 1), ref_count_used_list1=l1, is_ref=int(rng.random() < 0.7), qp=int(rng.choice([20, 35, 50, 63])),
              sc_class1=int(rng.random() < 0.15), rtc_tune=int(rng.random() < 0.15), similar_brightness_refs=int(rng.random() < 0.5),
              safe_limit_nref=int(rng.integers(0, 3)), only_l_bwd=int(rng.random() < 0.2), use_best_me_unipred_cand_only=int(rng.random() < 0.3),
              safe_limit_zz_th=int(rng.choice([0, 60000])), gm_level_enabled=int(rng.random() < 0.4),
              gm_use_distance_based_active_th=int(rng.random() < 0.5))
    nfr = 10
    motion = (int(rng.integers(-9, 10)), int(rng.integers(-6, 7)))
    seq = U.synth_sequence(W, H, nfr, seed=int(rng.integers(1, 1000)), motion=motion, noise=float(rng.choice([0.5, 2.0, 6.0])))
    if rng.random() < 0.15:
        seq[:] = seq[0]  # static clip
    if rng.random() < 0.1:
        seq = rng.integers(0, 256, size=seq.shape, dtype=np.uint8)
    cur = int(rng.integers(3, 7))
    past = sorted([n for n in range(nfr) if n < cur], reverse=True)
    fut = sorted([n for n in range(nfr) if n > cur])
    refs = {}
    for ri in range(l0):
        refs[(0, ri)] = past[ri % len(past)]
    for ri in range(l1):
        refs[(1, ri)] = (fut if rng.random() < 0.8 else past)[ri % 3]
    d = U.derive_input(preset, W, H, **kw)
    what = "preset %d %dx%d tl %d refs %s cur %d %s" % (preset, W, H, tl, refs, cur, {k: v for k, v in kw.items() if v})
    return d, seq, cur, refs, what


def params_for(d, cur, refs):
    """Parameters from the compiled reference where it is present (its own svt_aom_sig_deriv_me), else from the product library."""
    if U.ref_lib() is not None:
        return U.ref_params(d, cur, refs)
    import ctypes as C
    p = U.abi.SvtMeParams()
    assert U.product_lib().svtme_derive_params(C.byref(d), C.byref(p)) == 0
    return U.finish_params(p, d, cur, refs)


def cpu_result(which, p, seq, cur, refs):
    pyr = {n: U.CpuPyramid(which, seq[n], n, hme_level1=int(p.enable_hme_level1_flag)) for n in set(refs.values()) | {cur}}
    return U.cpu_me_picture(which, p, pyr[cur], {k: pyr[v] for k, v in refs.items()})


def differing(a, b):
    return [k for k in a if not np.array_equal(a[k], b[k])]


def run_cpu(seed, n):
    """oracle restatement vs compiled reference (C and AVX2 alternating); returns the descriptions of mismatching configurations."""
    rng, bad = np.random.default_rng(seed), []
    for it in range(n):
        d, seq, cur, refs, what = random_config(rng)
        try:
            p = params_for(d, cur, refs)
        except AssertionError:
            continue
        U.ref_lib().ref_set_simd(it & 1)
        diffs = differing(cpu_result("ref", p, seq, cur, refs), cpu_result("orc", p, seq, cur, refs))
        if diffs:
            bad.append((it, what, diffs))
    U.ref_lib().ref_set_simd(0)
    return bad


def run_gpu(seed, n):
    """CUDA path through the C ABI (parameters from svtme_derive_params) vs the CPU checker, deep state included."""
    import importlib
    me = importlib.import_module(U.PKG + ".me")
    rng, bad = np.random.default_rng(seed), []
    which = U.best_cpu_checker()
    for it in range(n):
        d, seq, cur, refs, what = random_config(rng)
        try:
            p = params_for(d, cur, refs)
        except AssertionError:
            continue
        want = cpu_result(which, p, seq, cur, refs)
        W, H = seq.shape[2], seq.shape[1]
        with me.MeSession(0, W, H, capacity=len(seq)) as ses:
            q = U.finish_params(ses.derive_params(d), d, cur, refs)
            if bytes(q) != bytes(p):
                bad.append((it, what, ["svtme_derive_params"]))
                continue
            need = sorted(set(refs.values()) | {cur})
            ses.put_pictures(need, [np.ascontiguousarray(seq[k]).ctypes.data for k in need], W)
            got = ses.run_picture(q, cur, refs, deep=True)
        diffs = differing(want, got)
        if diffs:
            bad.append((it, what, diffs))
    return bad
