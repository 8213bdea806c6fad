// svtme_common.cuh — device-side data contracts of the B200 open-loop ME path (sm_100a).
//
// HBM layout of one resident source picture ("pyramid", the device twin of EbPaReferenceObject,
// Source/Lib/Codec/reference_object.h:73-89): three 8-bit planes, each with replicate padding on
// all sides, rows 128-byte aligned so that a 64x64 block row / 16-byte search-window chunk is a
// single aligned vector load:
//     level 0 full      org (128, 72)  stride = up128(W   + 256)
//     level 1 quarter   org ( 64, 32)  stride = up128(W/2 + 128)
//     level 2 sixteenth org ( 32, 16)  stride = up128(W/4 +  64)
// The pads are at least the reference's (72/32/16, enc_handle.c:1276-1311,4082-4094); the LOGICAL
// search limits (63 / 31 / 15 px, motion_estimation.c:838-839,942-943,1043-1044,1257-1258) are
// parameters of the clipping code, not of this layout, so results do not depend on it.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>
#include "svtme_b200.h"

#define SVTME_NL SVTME_MAX_LISTS
#define SVTME_NR SVTME_MAX_REFS

struct DevPlane {
    uint8_t *base;   // first byte of the allocation (row 0 of the top pad)
    int stride;      // bytes per row
    int org_x, org_y;
    int width, height;
    __host__ __device__ const uint8_t *at(int x, int y) const {
        return base + (size_t)(org_y + y) * stride + org_x + x;
    }
};

struct DevPyramid {
    DevPlane lvl[3];
};

// One picture's ME job as the stage kernel sees it (array in device memory, one per picture).
struct DevJob {
    SvtMeParams prm;
    DevPyramid cur;
    DevPyramid ref[SVTME_NL][SVTME_NR];
    int width, height;      // aligned luma size
    int b64_cols, b64_rows; // ceil(width/64), ceil(height/64)
    int n_pu;               // exported PU count (85/21/5)
    // outputs (device), layouts of SvtMeOut
    uint8_t *total;
    uint32_t *mv;
    uint8_t *cand;
    uint32_t *dist64, *dist32, *dist16, *dist8, *var8, *rc_dist;
    uint8_t *stationary, *allow_gm;
    uint32_t *best_sad, *best_mv; // deep state, may be null
    int16_t *hme_sc;
    unsigned long long *hme_sad;
    uint8_t *do_ref;
    unsigned long long *work_ops; // [SVTME_PHASES], per launch slot: algorithmic SAD byte-ops executed for this picture (roofline numerator)
};

// launch slots of the stage pipeline (svtme_stage.cu), timed with one CUDA event after each
#define SVTME_PHASES 8
#define SVTME_PHASE_NAMES "state_clear", "zz_sad", "prehme", "hme_l0", "hme_l1", "hme_l2_centre", "integer_search", "finish"

static inline int up128(int v) { return (v + 127) & ~127; }

// kernels (svtme_pyramid.cu)
void svtme_launch_pad_plane(const DevPlane &pl, int pad_l, int pad_r, int pad_t, int pad_b, cudaStream_t st);
void svtme_launch_downsample(const DevPlane &full, const DevPlane &quarter, const DevPlane &sixteenth,
                             int hme_level1, cudaStream_t st);
void svtme_launch_downsample_generic(const uint8_t *in, int in_stride, int w, int h, uint8_t *out, int out_stride,
                                     int step, cudaStream_t st);
// batched pyramid build: 2 launches for n pictures
struct SvtmePutItem {
    const uint8_t *src; // device luma of the caller, 16-byte aligned rows; null = level 0 already holds the picture
    int src_stride;     // bytes
    int src_bits;       // 8: one byte per sample; 10: 16-bit samples, level 0 = sample >> 2 (svt_enc_msb_un_pack2_d, pack_unpack_c.c:202)
    DevPyramid pyr;
};
struct SvtmePadSpec {
    int l[3], r[3], t[3], b[3];
};
void svtme_launch_build_batch(const SvtmePutItem *d_items, int n, const DevPyramid &geom, const SvtmePadSpec &ps,
                              cudaStream_t st);
// transfers of the main stream that must not queue on a copy engine (svtme_pyramid.cu)
cudaError_t svtme_launch_fetch_from_host(void *dst, const void *src_pinned, size_t bytes, cudaStream_t st);
cudaError_t svtme_launch_zero_fill(void *dst, size_t bytes, cudaStream_t st);
// stage kernel (svtme_stage.cu)
// d_states: n_jobs * b64_per_job * svtme_stage_state_bytes() bytes of scratch; ev: SVTME_PHASES events or null
// What the pictures of a batch need, so that launches no picture uses are skipped.
struct SvtmeStageFlags {
    bool small_areas;     // every full-pel search area is at most svtme_fp_direct_max() positions
    bool zz, prehme, prehme_l1_round, hme_l0, hme_l0_round_b, hme_l1, hme_l2;
    bool int_dependent;   // some picture derives the other references' full-pel areas from (list 0, ref 0)'s result
    int max_refs[2];      // references searched per list, maximum over the batch
    bool l0_strip;        // HME level 0 in strip mode too (needs prehme_strip's condition; off: one CTA per b64)
    bool refine_small;    // HME level 1 / 2 windows are small in every picture: one warp per window, no staging
    bool prehme_strip;    // pre-HME runs in strip mode (one CTA per b64 column / row and reference): every
                          // picture of the batch uses the sub-sampled SAD (16x8 source block = one sub-block)
};
SvtmeStageFlags svtme_stage_flags(const SvtMeParams &p);
void svtme_configure_stage_kernels(); // per-device kernel attributes (dynamic shared memory), idempotent
// sev (optional): 4 * SVTME_PHASES events, [(2 * slot + round) * 2 + {begin, end}] around the search launches of a slot (the
// scalar bookkeeping launches that share the slot stay outside); *sev_mask: bit 2 * slot + round set for every bracket recorded
void svtme_launch_me_stage(const DevJob *d_jobs, int n_jobs, int b64_cols, int b64_rows, void *d_states, cudaStream_t st,
                           cudaEvent_t *ev, const SvtmeStageFlags &f, cudaEvent_t *sev = nullptr, uint32_t *sev_mask = nullptr);
int svtme_fp_direct_max();
size_t svtme_stage_state_bytes();
// kernel-test entry points (svtme_stage.cu)
void svtme_launch_sad_loop_test(const uint8_t *d_src, int src_stride, const uint8_t *d_ref, int ref_stride,
                                int block_height, int block_width, int ref_row_step_raw, int skip_line, int sa_w,
                                int sa_h, unsigned long long *d_key, cudaStream_t st);
void svtme_launch_fullpel_test(const uint8_t *d_src, int src_stride, const uint8_t *d_ref, int ref_stride, int ox,
                               int oy, int sa_w, int sa_h, int sub_sad, uint32_t *d_best_sad, uint32_t *d_best_mv,
                               cudaStream_t st);
cudaError_t svtme_launch_fullpel_batch(const uint8_t *d_src, int src_stride, const uint8_t *d_ref, int ref_stride,
                                       const SvtMeFullpelItem *d_items, int n, int sub_sad, uint32_t *d_best_sad, uint32_t *d_best_mv,
                                       cudaStream_t st);
// dynamic-GOP detector HME (svtme_stage.cu)
size_t svtme_dg_acc_bytes();
void svtme_dg_acc_unpack(const void *host, SvtMeDgMetrics *out);
cudaError_t svtme_launch_dg_hme(const DevPlane &src16, const DevPlane &ref16, int b64_cols, int b64_rows, int sa, void *acc, cudaStream_t st);
// picture statistics (svtme_stats.cu)
size_t svtme_stats_scratch_bytes(int n_b64);
size_t svtme_stats_item_bytes();
void svtme_stats_set_item(void *items_host, int i, const DevPyramid &pyr);
int svtme_launch_picture_statistics(const DevPyramid &geom, const void *items, int n, const SvtMeStatsParams &prm, uint8_t *scratch,
                                    size_t *off_hist, size_t *off_tail, cudaStream_t st);
size_t svtme_stats_tail_bytes();
void svtme_stats_tail_unpack(const void *tail_host, unsigned long long *avg_intensity16, unsigned long long *avg_luma, uint16_t *pic_avg_variance);
extern unsigned long long g_svtme_launches;
