#!/bin/bash
# build_encoder.sh — builds the reference ENCODER (C-only, stock cmake) with the dump / replay hooks of svtme_dump.c
# from a scratch copy of /root/reference, and leaves the binary in oracle/_ref/SvtAv1EncApp_dump (git-ignored; it
# travels to the GPU box with the snapshot). TEST INFRASTRUCTURE ONLY. Nothing of the reference is copied into the
# repository: the patch is five inserted calls (two around the ME of a b64, one in the dynamic-GOP detector, one after
# the picture statistics, one after the temporal filter's ME call), applied by pattern to the scratch copy.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REF:-/root/reference}"
WORK="${WORK:-/tmp/svtref_dump}"
OUT="$HERE/../_ref"
test -d "$REF/Source" || { echo "reference tree $REF not present"; exit 1; }
rm -rf "$WORK" && mkdir -p "$WORK/src" "$WORK/build" "$WORK/bin" "$OUT"
cp -r "$REF"/{CMakeLists.txt,Source,cmake,third_party,Config,test,gstreamer-plugin,ffmpeg_plugin} "$WORK/src/" 2>/dev/null || true
for f in LICENSE.md PATENTS.md README.md; do cp "$REF/$f" "$WORK/src/" 2>/dev/null || true; done
cp "$HERE/svtme_dump.c" "$HERE/../ref_ctx_map.h" "$HERE/../../include/svtme_b200.h" "$WORK/src/Source/Lib/Codec/"
python3 - "$WORK/src/Source/Lib/Codec" <<'PY'
import re, sys
d = sys.argv[1]
p = d + "/me_process.c"
s = open(p).read()
call = re.search(r"svt_aom_motion_estimation_b64\(pcs,\s*b64_index,\s*b64_origin_x,\s*b64_origin_y,\s*me_context_ptr->me_ctx,\s*input_pic\);", s)
assert call, "call site of svt_aom_motion_estimation_b64 not found in me_process.c"
hooked = ("svtme_hook_pre_b64(me_context_ptr->me_ctx);\n" + " " * 28 + call.group(0) + "\n" + " " * 28 +
          "if (in_results_ptr->task_type == TASK_PAME) svtme_hook_post_b64(pcs, b64_index, me_context_ptr->me_ctx, input_padded_pic);")
s = s[:call.start()] + hooked + s[call.end():]
decl = ("struct MeContext;\nvoid svtme_hook_pre_b64(struct MeContext *ctx);\n"
        "void svtme_hook_post_b64(PictureParentControlSet *pcs, uint32_t b64, struct MeContext *ctx, EbPictureBufferDesc *input_padded_pic);\n")
m = list(re.finditer(r'^#include .*$', s, re.M))[-1]
s = s[:m.end()] + "\n" + decl + s[m.end():]
open(p, "w").write(s)
# (2) the dynamic-GOP detector: metrics of a finished frame, taken under the detector's mutex right before the frame is
# signalled done (afterwards picture decision resets them for the next comparison)
p3 = d + "/pd_process.c"
s3 = open(p3).read()
dg = re.search(r"svt_post_semaphore\(ppcs->dg_detector->frame_done_sem\);", s3)
assert dg, "frame_done_sem post not found in pd_process.c"
s3 = s3[:dg.start()] + "{ svtme_hook_post_dg(ppcs); " + dg.group(0) + " }" + s3[dg.end():]
m3 = list(re.finditer(r'^#include .*$', s3, re.M))[-1]
s3 = s3[:m3.end()] + "\nvoid svtme_hook_post_dg(PictureParentControlSet *pcs);\n" + s3[m3.end():]
open(p3, "w").write(s3)
# (3) picture analysis: the statistics right after they are gathered
p2 = d + "/pic_analysis_process.c"
s2 = open(p2).read()
st = re.search(r"pa_ref_obj_->avg_luma\s*=\s*pcs->avg_luma;", s2)
assert st, "statistics call site not found in pic_analysis_process.c"
s2 = s2[:st.end()] + "\n                svtme_hook_post_statistics(scs, pcs);" + s2[st.end():]
m2 = list(re.finditer(r'^#include .*$', s2, re.M))[-1]
s2 = s2[:m2.end()] + "\nvoid svtme_hook_post_statistics(SequenceControlSet *scs, PictureParentControlSet *pcs);\n" + s2[m2.end():]
open(p2, "w").write(s2)
# (4) temporal filter: the ME call of produce_temporally_filtered_pic
p4 = d + "/temporal_filtering.c"
s4 = open(p4).read()
tf = re.search(r"svt_aom_motion_estimation_b64\(centre_pcs,.*?input_picture_ptr_central\);[^\n]*", s4, re.S)
assert tf, "TF call site of svt_aom_motion_estimation_b64 not found in temporal_filtering.c"
s4 = s4[:tf.end()] + ("\n                    svtme_hook_post_tf_b64(centre_pcs, (uint32_t)blk_row * blk_cols + blk_col, ctx, input_picture_ptr_central);"
                      ) + s4[tf.end():]
m4 = list(re.finditer(r'^#include .*$', s4, re.M))[-1]
s4 = s4[:m4.end()] + ("\nvoid svtme_hook_post_tf_b64(PictureParentControlSet *centre, uint32_t b64, struct MeContext *ctx, "
                      "EbPictureBufferDesc *central);\n") + s4[m4.end():]
open(p4, "w").write(s4)
c = d + "/CMakeLists.txt"
t = open(c).read()
assert "set(all_files" in t
t = t.replace("set(all_files", "set(all_files\n        svtme_dump.c", 1)
open(c, "w").write(t)
print("patched", p, "and", c)
PY
cd "$WORK/build"
cmake -G Ninja "$WORK/src" -DCMAKE_BUILD_TYPE=Release -DCOMPILE_C_ONLY=ON -DBUILD_TESTING=OFF -DBUILD_SHARED_LIBS=OFF \
      -DSVT_AV1_LTO=OFF -DBUILD_DEC=OFF -DCMAKE_OUTPUT_DIRECTORY="$WORK/bin" > "$WORK/cmake.log" 2>&1 || { tail -30 "$WORK/cmake.log"; exit 1; }
ninja SvtAv1EncApp > "$WORK/ninja.log" 2>&1 || { tail -40 "$WORK/ninja.log"; exit 1; }
cp "$WORK/bin/SvtAv1EncApp" "$OUT/SvtAv1EncApp_dump"
echo "built $OUT/SvtAv1EncApp_dump"
