"""CPU tests (-m "not gpu") of the host side: parameter derivation against the reference's own
svt_aom_sig_deriv_me (golden grid), ABI symbol export, struct layout, error behaviour."""
import ctypes as C
import gzip
import importlib
import json
import os
import re

import numpy as np
import pytest

import _util as U

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_library_exports_every_header_symbol():
    lib = U.product_lib()  # loads csrc/libsvtme_b200.so; no compute calls here
    hdr = open(os.path.join(U.ROOT, "include", "svtme_b200.h")).read()
    names = set(re.findall(r"\b(svtme_[a-z0-9_]+|svt_[a-z0-9_]+_b200(?:_[a-z0-9]+)?)\s*\(", hdr))
    assert "svt_sad_loop_kernel_b200_noalloc" in names
    assert len(names) >= 25
    for n in sorted(names):
        assert hasattr(lib, n), "symbol missing from libsvtme_b200.so: " + n


def test_struct_layout_matches_library():
    U.abi.check_layout(U.product_lib())
    assert U.oracle_lib().orc_sizeof_params() == C.sizeof(U.abi.SvtMeParams)
    assert U.oracle_lib().orc_sizeof_out() == C.sizeof(U.abi.SvtMeOut)


def test_derive_params_matches_reference_grid():
    lib = U.product_lib()
    grid = json.load(gzip.open(os.path.join(GOLD, "params.json.gz"), "rt"))
    assert len(grid) == 600
    skip = {"num_of_list_to_search", "num_of_ref_pic_to_search", "temporal_layer_index", "is_ref", "hierarchical_levels",
            "similar_brightness_refs", "max_number_of_pus_per_sb", "only_l_bwd", "gm_enabled",
            "gm_use_distance_based_active_th", "picture_number", "ref_picture_number"}  # per-picture, not from sig_deriv_me
    for row in grid:
        d = U.abi.SvtMeDeriveInput()
        for k, v in row["in"].items():
            setattr(d, k, v)
        p = U.abi.SvtMeParams()
        assert lib.svtme_derive_params(C.byref(d), C.byref(p)) == 0
        got = p.as_dict()
        for k, v in row["out"].items():
            if k in skip:
                continue
            assert got[k] == v, (row["in"], k, got[k], v)


def test_derive_params_rejects_bad_preset():
    lib = U.product_lib()
    d = U.derive_input(14, 1920, 1080)
    p = U.abi.SvtMeParams()
    assert lib.svtme_derive_params(C.byref(d), C.byref(p)) == -1  # SVTME_ERR_INVALID


def test_exported_pu_count():
    lib = U.product_lib()
    p = U.abi.SvtMeParams()
    p.max_number_of_pus_per_sb = 85
    for e16, e8, want in [(1, 1, 85), (1, 0, 21), (0, 0, 5)]:
        p.enable_me_16x16, p.enable_me_8x8 = e16, e8
        assert lib.svtme_exported_pu_count(C.byref(p)) == want == U.abi.exported_pu_count(p)


def test_missing_library_is_a_hard_error(monkeypatch):
    monkeypatch.setattr(U.abi, "_lib", None)
    monkeypatch.setattr(U.abi, "LIB_PATH", "/nonexistent/libsvtme_b200.so")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        U.abi.load_library()


def test_synth_sequence_is_deterministic():
    a = U.synth_sequence(128, 64, 3, seed=4)
    b = U.synth_sequence(128, 64, 3, seed=4)
    assert np.array_equal(a, b) and a.dtype == np.uint8 and a.shape == (3, 64, 128)


def test_pyramid_sharing_plan():
    """Picture p is built on rank p % N; a broadcast is planned for every picture another rank references,
    in one global order (every rank must issue the same collectives)."""
    shard = importlib.import_module(U.PKG + ".shard")
    gop = importlib.import_module(U.PKG + ".gop")
    jobs = gop.ra_jobs(33, 5, 3, 2)
    specs = [(None, j.picture_number, j.refs) for j in jobs]
    for world in (2, 4, 8):
        needed = [shard.pictures_for(specs, shard.rank_plan(len(specs), world, r)) for r in range(world)]
        plan = shard.sharing_plan(needed, world)
        assert [p for p, _, _ in plan] == sorted(p for p, _, _ in plan)
        for pic, owner, users in plan:
            assert owner == pic % world and owner not in users and users
            for r in users:
                assert pic in needed[r]
        planned = {p: set(u) for p, _, u in plan}
        for r in range(world):  # everything a rank needs is either its own build or comes by broadcast
            for pic in needed[r]:
                assert pic % world == r or r in planned[pic]


@pytest.mark.skipif(not os.path.isdir("/root/reference/Source"), reason="needs the reference headers (/root/reference)")
def test_integration_hooks_compile_against_the_reference_headers():
    """INTEGRATION.md quotes oracle/encoder_dump/svtme_hooks_example.c: it must compile (gcc -fsyntax-only) against the
    reference's own headers and include/svtme_b200.h."""
    import subprocess
    r = subprocess.run(["make", "-s", "-C", os.path.join(U.ROOT, "oracle"), "integration-check"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
