"""Reader / writer of the dump files of oracle/encoder_dump/svtme_dump.c (the REAL reference encoder with two hooks
around svt_aom_motion_estimation_b64). Test infrastructure only."""
import ctypes as C
import os

import numpy as np

import _util as U

abi = U.abi
HEADER_BYTES = 4096


class DumpHeader(C.Structure):
    _fields_ = [("magic", C.c_char * 8)] + [(n, C.c_uint32) for n in (
        "version", "header_bytes", "record_bytes", "width", "height", "b64_cols", "b64_rows", "n_pu", "max_refs", "max_cand", "max_l0",
        "deep", "input_resolution", "me_segments", "decode_order", "slice_type")] + [
        ("picture_number", C.c_uint64), ("ref_picture_number", (C.c_uint64 * 4) * 2),
        ("din", abi.SvtMeDeriveInput), ("params", abi.SvtMeParams)]


def up8(v):
    return (v + 7) & ~7


def record_layout(n_pu, max_refs, max_cand, deep):
    """[(name, dtype, count, byte offset)] of one b64 record + the record size."""
    items, o = [], 0
    for name, dt, cnt in (("total_me_candidate_index", "u1", n_pu), ("me_mv_array", "<u4", n_pu * max_refs),
                          ("me_candidate_array", "u1", n_pu * max_cand)):
        items.append((name, dt, cnt, o))
        o += up8(cnt * np.dtype(dt).itemsize)
    for k, name in enumerate(("me_64x64_distortion", "me_32x32_distortion", "me_16x16_distortion", "me_8x8_distortion",
                              "me_8x8_cost_variance", "rc_me_distortion")):
        items.append((name, "<u4", 1, o + 4 * k))
    items += [("stationary_block_present_sb", "u1", 1, o + 24), ("rc_me_allow_gm", "u1", 1, o + 25)]
    o += up8(26)
    if deep:
        for name, dt, cnt in (("p_sb_best_sad", "<u4", 680), ("p_sb_best_mv", "<u4", 680), ("hme_sc", "<i2", 16), ("hme_sad", "<u8", 8),
                              ("do_ref", "u1", 8)):
            items.append((name, dt, cnt, o))
            o += up8(cnt * np.dtype(dt).itemsize)
    return items, o


def read_dump(path):
    """-> (DumpHeader, {name: array [n_b64, count]})"""
    raw = open(path, "rb").read()
    h = DumpHeader.from_buffer_copy(raw[:C.sizeof(DumpHeader)])
    assert h.magic == b"SVTMEDMP" and h.header_bytes == HEADER_BYTES
    n_b64 = h.b64_cols * h.b64_rows
    items, rb = record_layout(h.n_pu, h.max_refs, h.max_cand, h.deep)
    assert rb == h.record_bytes and len(raw) == HEADER_BYTES + n_b64 * rb, (rb, h.record_bytes, len(raw))
    recs = np.frombuffer(raw, dtype=np.uint8, offset=HEADER_BYTES).reshape(n_b64, rb)
    out = {}
    for name, dt, cnt, o in items:
        a = np.ascontiguousarray(recs[:, o:o + cnt * np.dtype(dt).itemsize]).view(dt).reshape(n_b64, cnt)
        out[name] = a.reshape(-1) if cnt == 1 else a
    return h, out


def write_dump(path, header, arrays):
    """Same layout from result arrays (e.g. the GPU outputs) without deep state: input of SVTME_REPLAY_DIR."""
    h = DumpHeader.from_buffer_copy(bytes(header))
    h.deep = 0
    items, rb = record_layout(h.n_pu, h.max_refs, h.max_cand, 0)
    h.record_bytes = rb
    n_b64 = h.b64_cols * h.b64_rows
    recs = np.zeros((n_b64, rb), dtype=np.uint8)
    for name, dt, cnt, o in items:
        a = np.ascontiguousarray(arrays[name]).reshape(n_b64, cnt).astype(dt, copy=False)
        recs[:, o:o + cnt * np.dtype(dt).itemsize] = a.view(np.uint8).reshape(n_b64, -1)
    with open(path, "wb") as f:
        f.write(bytes(h).ljust(HEADER_BYTES, b"\0"))
        f.write(recs.tobytes())


def read_luma(dump_dir, pic, width, height):
    return np.fromfile(os.path.join(dump_dir, "luma_%06d.bin" % pic), dtype=np.uint8).reshape(height, width)


def struct_fields_equal(a, b, skip=()):
    """Field-by-field comparison of two ctypes structs of the same type; returns the names that differ."""
    bad = []
    for name, _ in a._fields_:
        if name.startswith("_pad") or name in skip:
            continue
        va, vb = getattr(a, name), getattr(b, name)
        if isinstance(va, (C.Structure, C.Array)):
            if bytes(va) != bytes(vb):
                bad.append(name)
        elif va != vb:
            bad.append(name)
    return bad


def read_stats(path):
    """stats_%06d.bin of oracle/encoder_dump/svtme_dump.c (svtme_hook_post_statistics): parameters and results of the
    encoder's own svt_aom_gathering_picture_statistics for one picture."""
    raw = open(path, "rb").read()
    hd = np.frombuffer(raw, dtype=np.uint32, count=16)
    assert hd[0] == 0x54535653
    n_b64, n_var = int(hd[3]), int(hd[4])
    off = 64
    out = {"params": hd[5:12].astype(np.int64), "size": hd[1:3].astype(np.int64),
           "pic_avg_variance": np.array([hd[12]], np.uint16)}
    out["avg_luma"] = np.frombuffer(raw, dtype=np.uint64, count=1, offset=off).copy(); off += 8
    out["average_intensity_per_region"] = np.frombuffer(raw, dtype=np.uint64, count=16, offset=off).reshape(4, 4).copy(); off += 128
    out["picture_histogram"] = np.frombuffer(raw, dtype=np.uint32, count=16 * 256, offset=off).reshape(4, 4, 256).copy(); off += 16 * 256 * 4
    if hd[6]:
        out["variance"] = np.frombuffer(raw, dtype=np.uint16, count=n_b64 * n_var, offset=off).reshape(n_b64, n_var).copy()
    return out


def read_dg(path):
    """dg_%06d_%06d.bin (svtme_hook_post_dg): (tot_dist, tot_cplx, tot_active, sum_in_vectors, input_resolution)."""
    raw = open(path, "rb").read()
    dist = int(np.frombuffer(raw, dtype=np.uint64, count=1)[0])
    r = np.frombuffer(raw, dtype=np.uint32, count=4, offset=8)
    return [dist, int(r[0]), int(r[1]), int(np.int32(r[2])), int(r[3])]


class TfDumpHeader(C.Structure):
    _fields_ = [("magic", C.c_char * 8)] + [(n, C.c_uint32) for n in ("width", "height", "b64_cols", "b64_rows", "record_bytes", "reserved")] + [
        ("centre", C.c_uint64), ("frame", C.c_uint64), ("tin", abi.SvtMeTfDeriveInput), ("params", abi.SvtMeParams)]


def read_tf_dump(path):
    """tf_%06d_%06d.bin (svtme_hook_post_tf_b64): header + per block the tables / HME results of (list 0, ref 0)."""
    raw = open(path, "rb").read()
    h = TfDumpHeader.from_buffer_copy(raw[:C.sizeof(TfDumpHeader)])
    assert h.magic == b"SVTMETF1" and h.record_bytes == 696
    n = h.b64_cols * h.b64_rows
    assert len(raw) == HEADER_BYTES + n * 696, (len(raw), n)
    recs = np.frombuffer(raw, dtype=np.uint8, offset=HEADER_BYTES).reshape(n, 696)
    return h, {"p_sb_best_sad": np.ascontiguousarray(recs[:, 0:340]).view("<u4"), "p_sb_best_mv": np.ascontiguousarray(recs[:, 340:680]).view("<u4"),
               "hme_sc": np.ascontiguousarray(recs[:, 680:684]).view("<i2"), "hme_sad": np.ascontiguousarray(recs[:, 688:696]).view("<u8").reshape(n)}
