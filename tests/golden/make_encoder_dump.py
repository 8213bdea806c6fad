"""Generates tests/golden/encdump_*.npz from the REAL reference encoder (run in the container, where /root/reference
exists): oracle/encoder_dump/build_encoder.sh builds SvtAv1EncApp (C-only, stock cmake) from a scratch copy of the
reference with two hook calls around svt_aom_motion_estimation_b64; this script encodes synthetic clips with
SVTME_DUMP_DIR set and packs, per ME picture, what the encoder handed to ME (SvtMeParams rebuilt from its MeContext,
the encoder-level SvtMeDeriveInput, the reference picture numbers) and what ME returned (MeSbResults arrays, per-b64
pcs arrays, optionally the deep state). The source luma is not stored: it is tests' synth_sequence(seed), and the
fixture keeps a CRC32 of every plane the encoder's ME actually read.

    python tests/golden/make_encoder_dump.py            # all cases
"""
import os
import subprocess
import sys
import tempfile
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import _encdump as E
import _util as U

APP = os.path.join(U.ROOT, "oracle", "_ref", "SvtAv1EncApp_dump")
# name, width, height, frames, preset, deep state, extra encoder arguments
CASES = [
    ("360p_p8", 640, 360, 16, 8, True, []),
    ("360p_p4", 640, 360, 10, 4, False, []),
    ("1080p_p8", 1920, 1080, 9, 8, False, []),
    # 10-bit source (16-bit samples = the 8-bit pattern << 2 plus two seeded LSBs, see samples16): ME reads the 8 MSBs
    ("360p_p6_10bit", 640, 360, 9, 6, False, ["--input-depth", "10"]),
    ("360p_p2", 640, 360, 6, 2, False, []),                         # slow preset: large areas, global-motion detection on
    ("360p_p11", 640, 360, 10, 11, False, []),                      # fast preset
    ("360p_p9_ld", 640, 360, 9, 9, False, ["--pred-struct", "1"]),  # low delay: list 0 only ... as picture decision posts it
]


def samples16(seq):
    """The 10-bit source of a '--input-depth 10' case: what bench.py's 10-bit workloads and the GPU tests upload."""
    rng = np.random.default_rng(3)
    return (seq.astype(np.uint16) << 2) | rng.integers(0, 4, size=seq.shape, dtype=np.uint16)
# aux-only case (no ME dump kept): long enough for picture decision to evaluate a 32-picture mini-GOP, which is when the
# dynamic-GOP detector compares pictures (eval_sub_mini_gop -> early_hme, pd_process.c:591-720)
AUX_ONLY = [("dg_360p_p8", 640, 360, 67, 8, False, [])]
# temporal filter on: the ME_MCTF calls of produce_temporally_filtered_pic (name, width, height, frames, preset)
TF_CASES = [("tf_360p_p8", 640, 360, 17, 8), ("tf_360p_p4", 640, 360, 9, 4)]
ENC_ARGS = ["--fps", "30", "--lp", "1", "--enable-tf", "0", "--scm", "0"]


def write_yuv(path, seq, bits=8):
    n, h, w = seq.shape
    chroma = (np.full((h // 2) * (w // 2) * 2, 128, np.uint8) if bits == 8 else np.full((h // 2) * (w // 2) * 2, 512, "<u2")).tobytes()
    src = seq if bits == 8 else samples16(seq).astype("<u2")
    with open(path, "wb") as f:
        for i in range(n):
            f.write(src[i].tobytes())
            f.write(chroma)


def encode(yuv, w, h, n, preset, extra, ivf, env):
    cmd = [APP, "-i", yuv, "-w", str(w), "-h", str(h), "-n", str(n), "--preset", str(preset)] + ENC_ARGS + extra + ["-b", ivf]
    subprocess.run(cmd, check=True, env=dict(os.environ, **env), stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)


def make_case(name, w, h, n, preset, deep, extra):
    seq = U.synth_sequence(w, h, n, seed=1)
    with tempfile.TemporaryDirectory() as tmp:
        yuv, ddir = os.path.join(tmp, "in.yuv"), os.path.join(tmp, "dump")
        os.mkdir(ddir)
        write_yuv(yuv, seq, 10 if "--input-depth" in extra else 8)
        env = {"SVTME_DUMP_DIR": ddir}
        if deep:
            env["SVTME_DUMP_DEEP"] = "1"
        encode(yuv, w, h, n, preset, extra, os.path.join(tmp, "out.ivf"), env)
        pack = {"meta": np.array([w, h, n, preset, int(deep)], dtype=np.int64)}
        pics = sorted(int(f[3:9]) for f in os.listdir(ddir) if f.startswith("me_"))
        pack["pictures"] = np.array(pics, dtype=np.int64)
        crc = {}
        for f in os.listdir(ddir):
            if f.startswith("luma_"):
                pic = int(f[5:11])
                plane = E.read_luma(ddir, pic, w, h)
                assert np.array_equal(plane, seq[pic]), "the encoder's ME input differs from the synthetic frame %d" % pic
                crc[pic] = zlib.crc32(plane.tobytes())
        pack["luma_pictures"] = np.array(sorted(crc), dtype=np.int64)
        pack["luma_crc32"] = np.array([crc[k] for k in sorted(crc)], dtype=np.uint32)
        for pic in pics:
            hdr, arrays = E.read_dump(os.path.join(ddir, "me_%06d.bin" % pic))
            pack["hdr_%d" % pic] = np.frombuffer(bytes(hdr), dtype=np.uint8)
            for k, a in arrays.items():
                pack["%s_%d" % (k, pic)] = a
        if (name, w, h, n, preset, deep, extra) not in AUX_ONLY:
            out = os.path.join(HERE, "encdump_%s.npz" % name)
            np.savez_compressed(out, **pack)
            print(name, "pictures", pics, "->", out, os.path.getsize(out) // 1024, "KiB")
        # what the same encode computed next to ME: picture statistics (every picture) and the dynamic-GOP detector's metrics
        aux = {"meta": np.array([w, h, n, preset], dtype=np.int64)}
        spics = sorted(int(f[6:12]) for f in os.listdir(ddir) if f.startswith("stats_"))
        if n > 20:
            spics = spics[:4]  # (the long clip is there for the DG detector)
        aux["stats_pictures"] = np.array(spics, dtype=np.int64)
        for pic in spics:
            for k, a in E.read_stats(os.path.join(ddir, "stats_%06d.bin" % pic)).items():
                aux["%s_%d" % (k, pic)] = a
        dgs = sorted(f for f in os.listdir(ddir) if f.startswith("dg_"))
        aux["dg"] = np.array([[int(f[3:9]), int(f[10:16])] + E.read_dg(os.path.join(ddir, f)) for f in dgs], dtype=np.int64).reshape(-1, 7)
        out = os.path.join(HERE, "encdump_aux_%s.npz" % name)
        np.savez_compressed(out, **aux)
        print(name, "statistics of", len(spics), "pictures,", len(dgs), "DG-detector frames ->", out, os.path.getsize(out) // 1024, "KiB")


def make_tf_case(name, w, h, n, preset):
    """encdump_tf_*.npz: per (centre picture, neighbouring frame) whose two planes are untouched synthetic frames, the
    SvtMeTfDeriveInput of the encoder, the SvtMeParams rebuilt from its MeContext, and the tables / HME results per block."""
    seq = U.synth_sequence(w, h, n, seed=1)
    with tempfile.TemporaryDirectory() as tmp:
        yuv, ddir = os.path.join(tmp, "in.yuv"), os.path.join(tmp, "dump")
        os.mkdir(ddir)
        write_yuv(yuv, seq)
        cmd = [APP, "-i", yuv, "-w", str(w), "-h", str(h), "-n", str(n), "--preset", str(preset), "--fps", "30", "--lp", "1",
               "--enable-tf", "1", "--scm", "0", "-b", os.path.join(tmp, "out.ivf")]
        subprocess.run(cmd, check=True, env=dict(os.environ, SVTME_DUMP_DIR=ddir), stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        pack, pairs, dropped = {"meta": np.array([w, h, n, preset], dtype=np.int64)}, [], 0
        for f in sorted(os.listdir(ddir)):
            if not f.startswith("tf_"):
                continue
            c, fr = int(f[3:9]), int(f[10:16])
            planes = [np.fromfile(os.path.join(ddir, "tfluma_%06d_%06d_%s.bin" % (c, fr, t)), dtype=np.uint8).reshape(h, w) for t in "cf"]
            if not (np.array_equal(planes[0], seq[c]) and np.array_equal(planes[1], seq[fr])):
                dropped += 1  # one of the two was already filtered: not reproducible from the synthetic clip
                continue
            hdr, arrays = E.read_tf_dump(os.path.join(ddir, f))
            pairs.append((c, fr))
            pack["hdr_%d_%d" % (c, fr)] = np.frombuffer(bytes(hdr), dtype=np.uint8)
            for k, a in arrays.items():
                pack["%s_%d_%d" % (k, c, fr)] = a
        pack["pairs"] = np.array(pairs, dtype=np.int64).reshape(-1, 2)
        out = os.path.join(HERE, "encdump_%s.npz" % name)
        np.savez_compressed(out, **pack)
        print(name, len(pairs), "(centre, frame) pairs kept,", dropped, "dropped ->", out, os.path.getsize(out) // 1024, "KiB")


if __name__ == "__main__":
    if not os.path.exists(APP):
        subprocess.check_call(["bash", os.path.join(U.ROOT, "oracle", "encoder_dump", "build_encoder.sh")])
    for case in CASES + AUX_ONLY:
        if len(sys.argv) < 2 or case[0] in sys.argv[1:]:
            make_case(*case)
    for case in TF_CASES:
        if len(sys.argv) < 2 or case[0] in sys.argv[1:]:
            make_tf_case(*case)
