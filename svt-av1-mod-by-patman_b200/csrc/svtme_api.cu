// svtme_api.cu — the C ABI of include/svtme_b200.h: picture-granular production entry points and
// the RTCD-signature kernel-test wrappers. Host code only orchestrates: every pixel operation is
// a CUDA kernel of this library; there is no CPU fallback (a CUDA failure is reported, never
// papered over).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <deque>
#include <map>
#include <mutex>
#include <vector>
#include "svtme_common.cuh"

static thread_local char g_err[512] = "";
static int failf(int code, const char *fmt, const char *who) {
    snprintf(g_err, sizeof(g_err), fmt, who);
    return code;
}
static int fail(int code, const char *what, cudaError_t e = cudaSuccess) {
    if (e != cudaSuccess)
        snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
    else
        snprintf(g_err, sizeof(g_err), "%s", what);
    return code;
}
#define CU(call)                                                  \
    do {                                                          \
        cudaError_t e_ = (call);                                  \
        if (e_ != cudaSuccess) return fail(SVTME_ERR_CUDA, #call, e_); \
    } while (0)

extern "C" const char *svtme_last_error(void) { return g_err; }
void svtme_set_error(const char *what) { snprintf(g_err, sizeof(g_err), "%s", what); } // for the other translation units
extern "C" uint64_t svtme_launch_count(void) { return g_svtme_launches; }
extern "C" uint32_t svtme_sizeof_params(void) { return (uint32_t)sizeof(SvtMeParams); }
extern "C" uint32_t svtme_sizeof_out(void) { return (uint32_t)sizeof(SvtMeOut); }
extern "C" uint32_t svtme_sizeof_job(void) { return (uint32_t)sizeof(SvtMeJob); }
extern "C" uint32_t svtme_sizeof_derive_input(void) { return (uint32_t)sizeof(SvtMeDeriveInput); }

#define SVTME_LAUNCH_EV_RING 64
#define SVTME_NS 4 // compute streams at most (the event ring is indexed by launch sequence number, seq % SVTME_NS = stream)
struct Slot {
    DevPyramid pyr;
    uint64_t pic_num;
    bool used;
    uint64_t last_reader[SVTME_NS]; // per compute stream: sequence number of the newest launch that reads this pyramid (a writer of the slot waits for both); 0 = none
};

static inline size_t up16(size_t v) { return (v + 15) & ~(size_t)15; }

struct OutBuffers { // device + pinned host staging for one batch of jobs
    size_t bytes = 0;
    uint8_t *dev = nullptr;
    uint8_t *host = nullptr;
};

// One launch of the stage pipeline: its job array, result buffers and events. The handle owns one
// Batch for the synchronous launch/fetch pair and a pool of them for the queued (pipelined) form,
// where several launches are in flight and results come back in submission order.
struct Batch {
    DevJob *d_jobs = nullptr;
    DevJob *h_jobs = nullptr; // pinned
    uint32_t jobs_cap = 0;
    OutBuffers ob;
    std::vector<size_t> work_off; // offsets of the per-job work counters
    std::vector<uint8_t> direct;  // queued form: job i's three big arrays were downloaded straight into the caller's page-locked arrays
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, jobs_ev = nullptr, done_ev = nullptr; // done_ev: download complete
    cudaEvent_t ph_ev[SVTME_PHASES] = {};
    cudaEvent_t sev[4 * SVTME_PHASES] = {}; // svtme_set_search_timing: brackets around the search launches of each slot (created on first use)
    uint32_t sev_mask = 0;
    bool jobs_pending = false; // h_jobs may still be on its way to the device
    uint32_t n = 0;
    size_t total = 0;
    uint64_t seq = 0;
};

#define SVTME_PUT_RING 4
struct PutRing { // descriptors of one batched pyramid build (device, pinned) + "copied" event
    SvtmePutItem *d = nullptr, *h = nullptr;
    uint32_t cap = 0;
    bool pending = false;
    cudaEvent_t ev = nullptr;
};

struct SvtMeHandle {
    int device;
    uint32_t width, height, b64_cols, b64_rows, n_b64;
    cudaStream_t stream;
    bool own_stream;
    cudaStream_t h2d_stream, d2h_stream; // uploads + pyramid builds / result downloads overlap the pipeline of other batches
    cudaStream_t xstream[SVTME_NS];      // extra compute streams ([0] unused = `stream`): queued batches rotate over n_streams
                                         // of them, so that the launch tails of one batch are filled by the next ones
    int n_streams = 2;
    cudaEvent_t origin_ev = nullptr; // svtme_debug_mark
    bool search_timing = false;      // svtme_set_search_timing
    uint8_t *stats_dev = nullptr;    // svtme_picture_statistics: device scratch and its pinned mirror (created on first use)
    uint8_t *stats_host = nullptr;
    size_t stats_cap = 0;
    void *dg_dev = nullptr, *dg_host = nullptr; // svtme_dg_detector_hme: metric accumulators and their pinned mirror
    cudaEvent_t stats_ev[3] = {}; // before the kernels, after them, after the download (svtme_last_statistics_ms)
    float span[3] = {0.f, 0.f, 0.f};  // last collected batch: pipeline start / end / download end, ms after the mark
    uint32_t submit_count = 0;
    void *x_states[SVTME_NS] = {};      // per extra compute stream: its own per-b64 scratch
    size_t x_states_bytes[SVTME_NS] = {};
    cudaEvent_t h2d_ev, build_ev;
    cudaEvent_t ext_ev = nullptr;   // svtme_external_write_end / svtme_share_picture: "externally written pyramids are complete"
    bool ext_recorded = false;
    cudaEvent_t launch_ev[SVTME_LAUNCH_EV_RING]; // "launch seq is done", indexed seq % ring (Batch objects and their events are recycled)
    bool build_recorded = false;
    uint8_t *u16_stage = nullptr; // 16-bit input pictures of the batched build in flight (svtme_put_pictures_10bit)
    size_t u16_bytes = 0;
    uint8_t *pool; // all planes of all slots
    size_t plane_bytes[3], slot_bytes;
    std::vector<Slot> slots;
    std::map<uint64_t, int> index;
    Batch *sync_batch = nullptr;   // svtme_launch_pictures / svtme_fetch_results
    Batch *last = nullptr;         // newest launch (measurement getters)
    std::vector<Batch *> free_batches;
    std::deque<Batch *> queue;     // submitted, not yet collected (FIFO)
    uint64_t seq = 0;
    size_t max_ob_bytes = 0;   // largest result record / job count any batch has needed so far: batch buffers are (re)sized
    uint32_t max_jobs = 0;     // to these, so that a steady stream of batches of mixed sizes stops reallocating early
    void *d_states = nullptr; // per-b64 scratch of the stage pipeline
    PutRing put[SVTME_PUT_RING];
    uint32_t put_seq = 0;
    size_t states_bytes = 0;
    float last_ms;
    std::mutex mu;
};

static Batch *batch_new() {
    Batch *b = new (std::nothrow) Batch();
    if (!b) return nullptr;
    cudaError_t e = cudaEventCreate(&b->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&b->ev1);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&b->jobs_ev, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreate(&b->done_ev);
    for (int k = 0; k < SVTME_PHASES && e == cudaSuccess; ++k) e = cudaEventCreate(&b->ph_ev[k]);
    if (e != cudaSuccess) { delete b; return nullptr; }
    return b;
}
static void batch_free(Batch *b) {
    if (!b) return;
    cudaFree(b->d_jobs);
    cudaFreeHost(b->h_jobs);
    cudaFree(b->ob.dev);
    cudaFreeHost(b->ob.host);
    if (b->ev0) cudaEventDestroy(b->ev0);
    if (b->ev1) cudaEventDestroy(b->ev1);
    if (b->jobs_ev) cudaEventDestroy(b->jobs_ev);
    if (b->done_ev) cudaEventDestroy(b->done_ev);
    for (int k = 0; k < SVTME_PHASES; ++k)
        if (b->ph_ev[k]) cudaEventDestroy(b->ph_ev[k]);
    for (int k = 0; k < 4 * SVTME_PHASES; ++k)
        if (b->sev[k]) cudaEventDestroy(b->sev[k]);
    delete b;
}

static const int kOrgX[3] = {128, 64, 32}, kOrgY[3] = {72, 32, 16};
// pads that are actually filled (>= the reference's 72/32/16 and >= every over-read of the kernels)
static const int kPadL[3] = {128, 64, 32}, kPadR[3] = {128, 64, 32}, kPadT[3] = {72, 32, 16}, kPadB[3] = {72, 32, 16};

static void layout_plane(DevPlane &pl, uint8_t *base, int level, int w, int h) {
    pl.base = base;
    pl.org_x = kOrgX[level];
    pl.org_y = kOrgY[level];
    pl.width = w >> level;
    pl.height = h >> level;
    pl.stride = up128(pl.width + kPadL[level] + kPadR[level]);
}
static size_t plane_size(int level, int w, int h) {
    DevPlane pl;
    layout_plane(pl, nullptr, level, w, h);
    // one extra row of slack: staging may touch the 16 bytes after the last pad row's last chunk
    return (size_t)pl.stride * (pl.height + kPadT[level] + kPadB[level] + 1);
}

extern "C" int svtme_open(int device, uint32_t width, uint32_t height, uint32_t capacity, SvtMeHandle **out) {
    if (!out || width < 64 || height < 64 || (width & 7) || (height & 7) || capacity == 0 || width > 16384 || height > 8704)
        return fail(SVTME_ERR_INVALID, "svtme_open: width/height must be multiples of 8, >= 64 (aligned luma size)");
    CU(cudaSetDevice(device));
    SvtMeHandle *h = new (std::nothrow) SvtMeHandle();
    if (!h) return fail(SVTME_ERR_NOMEM, "svtme_open: host allocation");
    h->device = device;
    h->width = width;
    h->height = height;
    h->b64_cols = (width + 63) / 64;
    h->b64_rows = (height + 63) / 64;
    h->n_b64 = h->b64_cols * h->b64_rows;
    h->slot_bytes = 0;
    for (int l = 0; l < 3; ++l) {
        h->plane_bytes[l] = (plane_size(l, width, height) + 255) & ~(size_t)255;
        h->slot_bytes += h->plane_bytes[l];
    }
    h->pool = nullptr;
    h->last_ms = 0.f;
    h->h2d_stream = h->d2h_stream = nullptr;
    for (int k = 0; k < SVTME_NS; ++k) h->xstream[k] = nullptr;
    {
        const char *ev = getenv("SVTME_COMPUTE_STREAMS");
        const int v = ev ? atoi(ev) : 2;
        h->n_streams = getenv("SVTME_SINGLE_STREAM") ? 1 : (v >= 1 && v <= SVTME_NS ? v : 2);
    }
    h->h2d_ev = h->build_ev = nullptr;
    cudaError_t e = cudaMalloc(&h->pool, h->slot_bytes * capacity);
    if (e != cudaSuccess) { delete h; return fail(SVTME_ERR_NOMEM, "svtme_open: cudaMalloc pyramid pool", e); }
    e = cudaMemset(h->pool, 0, h->slot_bytes * capacity);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    h->own_stream = true;
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->h2d_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->d2h_stream, cudaStreamNonBlocking);
    for (int k = 1; k < SVTME_NS && e == cudaSuccess; ++k) e = cudaStreamCreateWithFlags(&h->xstream[k], cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->h2d_ev, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->build_ev, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ext_ev, cudaEventDisableTiming);
    for (int k = 0; k < SVTME_PUT_RING && e == cudaSuccess; ++k) e = cudaEventCreateWithFlags(&h->put[k].ev, cudaEventDisableTiming);
    for (int k = 0; k < SVTME_LAUNCH_EV_RING; ++k) h->launch_ev[k] = nullptr;
    for (int k = 0; k < SVTME_LAUNCH_EV_RING && e == cudaSuccess; ++k) e = cudaEventCreateWithFlags(&h->launch_ev[k], cudaEventDisableTiming);
    if (e == cudaSuccess && !(h->sync_batch = batch_new())) e = cudaErrorMemoryAllocation;
    if (e != cudaSuccess) { cudaFree(h->pool); delete h; return fail(SVTME_ERR_CUDA, "svtme_open", e); }
    h->slots.resize(capacity);
    for (uint32_t i = 0; i < capacity; ++i) {
        uint8_t *b = h->pool + h->slot_bytes * i;
        for (int l = 0; l < 3; ++l) {
            layout_plane(h->slots[i].pyr.lvl[l], b, l, width, height);
            b += h->plane_bytes[l];
        }
        h->slots[i].used = false;
        h->slots[i].pic_num = 0;
        for (int k = 0; k < SVTME_NS; ++k) h->slots[i].last_reader[k] = 0;
    }
    *out = h;
    return SVTME_OK;
}

extern "C" void svtme_close(SvtMeHandle *h) {
    if (!h) return;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    for (int k = 1; k < SVTME_NS; ++k) cudaStreamSynchronize(h->xstream[k]);
    cudaStreamSynchronize(h->h2d_stream);
    cudaStreamSynchronize(h->d2h_stream);
    cudaFree(h->pool);
    for (int k = 1; k < SVTME_NS; ++k) {
        cudaFree(h->x_states[k]);
        cudaStreamDestroy(h->xstream[k]);
    }
    batch_free(h->sync_batch);
    for (Batch *b : h->free_batches) batch_free(b);
    for (Batch *b : h->queue) batch_free(b);
    cudaFree(h->d_states);
    cudaFree(h->u16_stage);
    for (int k = 0; k < SVTME_PUT_RING; ++k) {
        cudaFree(h->put[k].d);
        cudaFreeHost(h->put[k].h);
        cudaEventDestroy(h->put[k].ev);
    }
    cudaEventDestroy(h->h2d_ev);
    cudaEventDestroy(h->build_ev);
    if (h->ext_ev) cudaEventDestroy(h->ext_ev);
    if (h->origin_ev) cudaEventDestroy(h->origin_ev);
    cudaFree(h->stats_dev);
    cudaFreeHost(h->stats_host);
    cudaFree(h->dg_dev);
    cudaFreeHost(h->dg_host);
    for (cudaEvent_t e : h->stats_ev)
        if (e) cudaEventDestroy(e);
    for (int k = 0; k < SVTME_LAUNCH_EV_RING; ++k) cudaEventDestroy(h->launch_ev[k]);
    cudaStreamDestroy(h->h2d_stream);
    cudaStreamDestroy(h->d2h_stream);
    if (h->own_stream) cudaStreamDestroy(h->stream);
    delete h;
}

extern "C" uint32_t svtme_b64_count(const SvtMeHandle *h) { return h->n_b64; }

static int find_slot(SvtMeHandle *h, uint64_t pic, bool create) {
    auto it = h->index.find(pic);
    if (it != h->index.end()) return it->second;
    if (!create) return -1;
    for (size_t i = 0; i < h->slots.size(); ++i)
        if (!h->slots[i].used) {
            h->slots[i].used = true;
            h->slots[i].pic_num = pic;
            h->index[pic] = (int)i;
            return (int)i;
        }
    return -2;
}

// n pictures -> pyramids: (host input) n H2D copies, then 2 launches for the whole batch.
// bits = 8: `stride` bytes between rows of one-byte samples. bits = 10: 16-bit samples (10 significant bits), `stride` in
// SAMPLES; the planes get (uint8_t)(sample >> 2), what svt_aom_un_pack2d (pic_operators.c:228-244) leaves in the 8-bit
// buffer that ME reads.
static int put_batch(SvtMeHandle *h, uint32_t n, const uint64_t *pic_nums, const uint8_t *const *luma, uint32_t stride,
                     bool device_src, const char *who, int bits = 8) {
    if (!h || !pic_nums || !luma || stride < h->width) return failf(SVTME_ERR_INVALID, "%s: bad argument", who);
    const uint32_t stride_bytes = bits == 10 ? stride * 2u : stride;
    if (n == 0) return SVTME_OK;
    std::lock_guard<std::mutex> lk(h->mu);
    CU(cudaSetDevice(h->device));
    PutRing &pr = h->put[h->put_seq++ % SVTME_PUT_RING];
    if (pr.pending) CU(cudaEventSynchronize(pr.ev)); // this ring entry's descriptors have left the pinned buffer
    pr.pending = false;
    if (n > pr.cap) {
        cudaFree(pr.d);
        cudaFreeHost(pr.h);
        pr.d = pr.h = nullptr;
        pr.cap = 0;
        CU(cudaMalloc(reinterpret_cast<void **>(&pr.d), up16(sizeof(SvtmePutItem) * n)));
        CU(cudaMallocHost(reinterpret_cast<void **>(&pr.h), up16(sizeof(SvtmePutItem) * n)));
        pr.cap = n;
    }
    SvtmePutItem *items = pr.h;
    // Transactional: everything that can be rejected is checked before the ring is touched (a failure
    // at picture i > 0 used to leave pictures 0..i-1 indexed with no upload or build enqueued).
    {
        size_t fresh = 0, free_slots = 0;
        for (uint32_t i = 0; i < n; ++i) {
            if (!luma[i]) return failf(SVTME_ERR_INVALID, "%s: null picture", who);
            bool seen = h->index.count(pic_nums[i]) != 0;
            for (uint32_t k = 0; k < i && !seen; ++k) seen = pic_nums[k] == pic_nums[i];
            fresh += seen ? 0 : 1;
        }
        for (const Slot &sl : h->slots) free_slots += sl.used ? 0 : 1;
        if (fresh > free_slots) return failf(SVTME_ERR_NOMEM, "%s: pyramid ring full (release pictures or raise capacity)", who);
    }
    std::vector<int> created, touched;
    // a CUDA failure further down leaves the ring as it was: new index entries are dropped again
    struct Rollback {
        SvtMeHandle *h;
        std::vector<int> *created;
        bool armed = true;
        ~Rollback() {
            if (!armed) return;
            for (int si : *created) {
                h->index.erase(h->slots[si].pic_num);
                h->slots[si].used = false;
            }
        }
    } rollback{h, &created};
    // Host pictures travel on the upload stream so that they overlap the ME pipeline of earlier
    // batches; an upload into a slot must not overtake a launch that still reads the slot's old
    // content (all launches are in order on the main stream: waiting for the newest reader suffices).
    uint64_t newest[SVTME_NS] = {};
    bool any_copy = false;
    for (uint32_t i = 0; i < n; ++i) {
        const bool existed = h->index.count(pic_nums[i]) != 0;
        const int si = find_slot(h, pic_nums[i], true);
        if (si < 0) return failf(SVTME_ERR_NOMEM, "%s: pyramid ring full (release pictures or raise capacity)", who);
        if (!existed) created.push_back(si);
        touched.push_back(si);
        Slot &s = h->slots[si];
        for (int k = 0; k < SVTME_NS; ++k)
            if (s.last_reader[k] > newest[k]) newest[k] = s.last_reader[k]; // (cleared once the waits below are enqueued)
        items[i].pyr = s.pyr;
        items[i].src = nullptr;
        items[i].src_stride = 0;
        items[i].src_bits = bits;
        if (device_src && ((uintptr_t)luma[i] & 15) == 0 && (stride_bytes & 15) == 0) {
            items[i].src = luma[i]; // the decimation kernel reads the caller's buffer and fills level 0 itself
            items[i].src_stride = (int)stride_bytes;
        } else
            any_copy = true;
    }
    if (bits == 10 && any_copy) { // 16-bit pictures are copied to a staging area the decimation kernel converts from
        const size_t need = (size_t)n * h->width * 2 * h->height;
        if (need > h->u16_bytes) {
            CU(cudaStreamSynchronize(h->h2d_stream)); // an earlier build may still read the old staging area
            cudaFree(h->u16_stage);
            h->u16_stage = nullptr;
            h->u16_bytes = 0;
            CU(cudaMalloc(&h->u16_stage, need));
            h->u16_bytes = need;
        }
    }
    // Uploads AND pyramid builds run on the upload stream: a build never queues behind the ME pipeline of an
    // earlier batch. The stream first waits for the newest reader of the slots on each compute stream (an
    // event of the ring that has been re-recorded since belongs to a LATER launch of the same stream: still
    // sufficient); launches wait for build_ev.
    cudaStream_t cs = h->h2d_stream;
    if (device_src) { // the caller produced the pictures on the handle's stream (svtme_set_stream): keep that order
        CU(cudaEventRecord(h->h2d_ev, h->stream));
        CU(cudaStreamWaitEvent(cs, h->h2d_ev, 0));
    }
    for (int k = 0; k < SVTME_NS; ++k)
        if (newest[k]) CU(cudaStreamWaitEvent(cs, h->launch_ev[newest[k] % SVTME_LAUNCH_EV_RING], 0));
    if (h->ext_recorded) CU(cudaStreamWaitEvent(cs, h->ext_ev, 0));   // an exchange that reads or writes ring slots is still in flight
    for (int si : touched) // the writes below are ordered behind every reader now: the slots start a new life
        for (int k = 0; k < SVTME_NS; ++k) h->slots[si].last_reader[k] = 0;
    if (any_copy) {
        for (uint32_t i = 0; i < n; ++i) {
            if (items[i].src) continue;
            const DevPlane &f = items[i].pyr.lvl[0];
            // a pageable host buffer makes this copy synchronous w.r.t. the host
            if (bits == 10) { // (in-order on this stream after the previous batch's conversion: the staging area is free)
                uint8_t *dst = h->u16_stage + (size_t)i * h->width * 2 * h->height;
                CU(cudaMemcpy2DAsync(dst, (size_t)h->width * 2, luma[i], stride_bytes, (size_t)h->width * 2, h->height,
                                     device_src ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, cs));
                items[i].src = dst;
                items[i].src_stride = (int)h->width * 2;
            } else
                CU(cudaMemcpy2DAsync(const_cast<uint8_t *>(f.at(0, 0)), f.stride, luma[i], stride, h->width, h->height,
                                     device_src ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, cs));
        }
    }
    CU(svtme_launch_fetch_from_host(pr.d, items, up16(sizeof(SvtmePutItem) * n), cs));
    CU(cudaEventRecord(pr.ev, cs));
    pr.pending = true;
    SvtmePadSpec ps;
    for (int l = 0; l < 3; ++l) {
        ps.l[l] = kPadL[l]; ps.r[l] = kPadR[l]; ps.t[l] = kPadT[l]; ps.b[l] = kPadB[l];
    }
    svtme_launch_build_batch(pr.d, (int)n, items[0].pyr, ps, cs);
    CU(cudaGetLastError());
    CU(cudaEventRecord(h->build_ev, cs));
    h->build_recorded = true;
    rollback.armed = false;
    return SVTME_OK;
}

extern "C" int svtme_put_pictures(SvtMeHandle *h, uint32_t n, const uint64_t *pic_nums, const uint8_t *const *luma,
                                  uint32_t stride) {
    return put_batch(h, n, pic_nums, luma, stride, false, "svtme_put_pictures");
}

extern "C" int svtme_put_pictures_device(SvtMeHandle *h, uint32_t n, const uint64_t *pic_nums, const void *const *d_luma,
                                         uint32_t stride) {
    return put_batch(h, n, pic_nums, reinterpret_cast<const uint8_t *const *>(d_luma), stride, true,
                     "svtme_put_pictures_device");
}

extern "C" int svtme_put_pictures_10bit(SvtMeHandle *h, uint32_t n, const uint64_t *pic_nums, const uint16_t *const *luma16,
                                        uint32_t stride_samples) {
    return put_batch(h, n, pic_nums, reinterpret_cast<const uint8_t *const *>(luma16), stride_samples, false,
                     "svtme_put_pictures_10bit", 10);
}

extern "C" int svtme_put_pictures_10bit_device(SvtMeHandle *h, uint32_t n, const uint64_t *pic_nums, const void *const *d_luma16,
                                               uint32_t stride_samples) {
    return put_batch(h, n, pic_nums, reinterpret_cast<const uint8_t *const *>(d_luma16), stride_samples, true,
                     "svtme_put_pictures_10bit_device", 10);
}

extern "C" int svtme_put_picture(SvtMeHandle *h, uint64_t pic_num, const uint8_t *luma, uint32_t stride) {
    return put_batch(h, 1, &pic_num, &luma, stride, false, "svtme_put_picture");
}

extern "C" int svtme_put_picture_device(SvtMeHandle *h, uint64_t pic_num, const void *d_luma, uint32_t stride) {
    const uint8_t *p = reinterpret_cast<const uint8_t *>(d_luma);
    return put_batch(h, 1, &pic_num, &p, stride, true, "svtme_put_picture_device");
}

extern "C" int svtme_adopt_picture(SvtMeHandle *h, uint64_t pic_num) {
    if (!h) return fail(SVTME_ERR_INVALID, "svtme_adopt_picture: null handle");
    std::lock_guard<std::mutex> lk(h->mu);
    const int si = find_slot(h, pic_num, true);
    if (si == -2) return fail(SVTME_ERR_NOMEM, "svtme_adopt_picture: pyramid ring full");
    return SVTME_OK;
}

// ---- pyramids written by someone else (NCCL receive on the caller's stream, peer copy): ordering without host syncs.
// svtme_external_write_begin: the pictures get ring slots if they have none, and `cuda_stream` is made to wait for
// (a) the pyramid builds enqueued so far (a sender reads what svtme_put_pictures produced) and (b) every launch in
// flight that still reads one of those slots (a receiver overwrites them). svtme_external_write_end: everything
// enqueued on `cuda_stream` up to here must complete before any later launch, build or upload of this handle runs.
extern "C" int svtme_external_write_begin(SvtMeHandle *h, uint32_t n, const uint64_t *pic_nums, void *cuda_stream) {
    if (!h || (n && !pic_nums)) return fail(SVTME_ERR_INVALID, "svtme_external_write_begin: bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    CU(cudaSetDevice(h->device));
    cudaStream_t cs = (cudaStream_t)cuda_stream;
    {
        size_t fresh = 0, free_slots = 0;
        for (uint32_t i = 0; i < n; ++i) {
            bool seen = h->index.count(pic_nums[i]) != 0;
            for (uint32_t k = 0; k < i && !seen; ++k) seen = pic_nums[k] == pic_nums[i];
            fresh += seen ? 0 : 1;
        }
        for (const Slot &sl : h->slots) free_slots += sl.used ? 0 : 1;
        if (fresh > free_slots) return fail(SVTME_ERR_NOMEM, "svtme_external_write_begin: pyramid ring full");
    }
    uint64_t newest[SVTME_NS] = {};
    std::vector<int> touched;
    for (uint32_t i = 0; i < n; ++i) {
        const int si = find_slot(h, pic_nums[i], true);
        if (si < 0) return fail(SVTME_ERR_NOMEM, "svtme_external_write_begin: pyramid ring full");
        touched.push_back(si);
        for (int k = 0; k < SVTME_NS; ++k)
            if (h->slots[si].last_reader[k] > newest[k]) newest[k] = h->slots[si].last_reader[k];
    }
    if (h->build_recorded) CU(cudaStreamWaitEvent(cs, h->build_ev, 0));
    for (int k = 0; k < SVTME_NS; ++k)
        if (newest[k]) CU(cudaStreamWaitEvent(cs, h->launch_ev[newest[k] % SVTME_LAUNCH_EV_RING], 0));
    for (int si : touched)
        for (int k = 0; k < SVTME_NS; ++k) h->slots[si].last_reader[k] = 0;
    return SVTME_OK;
}

extern "C" int svtme_external_write_end(SvtMeHandle *h, void *cuda_stream) {
    if (!h) return fail(SVTME_ERR_INVALID, "svtme_external_write_end: null handle");
    std::lock_guard<std::mutex> lk(h->mu);
    CU(cudaSetDevice(h->device));
    CU(cudaEventRecord(h->ext_ev, (cudaStream_t)cuda_stream));
    h->ext_recorded = true;
    return SVTME_OK;
}

// One process, several GPUs (the reference is ONE process with a pool of ME threads): copy the pyramid of `pic_num`
// from the handle that built it into a ring slot of a handle on another device — cudaMemcpyPeerAsync over NVLink on
// the destination's upload stream, ordered after the source's builds and after the destination's launches that still
// read the slot; the destination's later launches wait for the copy, the source's next upload into the slot too.
extern "C" int svtme_share_picture(SvtMeHandle *src, SvtMeHandle *dst, uint64_t pic_num) {
    if (!src || !dst || src == dst) return fail(SVTME_ERR_INVALID, "svtme_share_picture: two different handles are needed");
    if (src->width != dst->width || src->height != dst->height || src->slot_bytes != dst->slot_bytes)
        return fail(SVTME_ERR_INVALID, "svtme_share_picture: the handles hold pictures of different sizes");
    std::mutex &m1 = src < dst ? src->mu : dst->mu, &m2 = src < dst ? dst->mu : src->mu; // fixed lock order
    std::lock_guard<std::mutex> l1(m1), l2(m2);
    const int ss = find_slot(src, pic_num, false);
    if (ss < 0) return fail(SVTME_ERR_NOTFOUND, "svtme_share_picture: picture not resident on the source handle");
    if (dst->index.count(pic_num) == 0) {
        bool any_free = false;
        for (const Slot &sl : dst->slots) any_free |= !sl.used;
        if (!any_free) return fail(SVTME_ERR_NOMEM, "svtme_share_picture: destination pyramid ring full");
    }
    CU(cudaSetDevice(dst->device));
    const int ds = find_slot(dst, pic_num, true);
    cudaStream_t cs = dst->h2d_stream;
    if (src->build_recorded) CU(cudaStreamWaitEvent(cs, src->build_ev, 0));
    if (src->ext_recorded) CU(cudaStreamWaitEvent(cs, src->ext_ev, 0));
    Slot &d = dst->slots[ds];
    for (int k = 0; k < SVTME_NS; ++k) {
        if (d.last_reader[k]) CU(cudaStreamWaitEvent(cs, dst->launch_ev[d.last_reader[k] % SVTME_LAUNCH_EV_RING], 0));
        d.last_reader[k] = 0;
    }
    CU(cudaMemcpyPeerAsync(d.pyr.lvl[0].base, dst->device, src->slots[ss].pyr.lvl[0].base, src->device, src->slot_bytes, cs));
    CU(cudaEventRecord(dst->build_ev, cs));
    dst->build_recorded = true;
    // the copy READS the source's slot: the source's upload stream (in order with its later uploads) waits for it. (An
    // event can only be recorded on a stream of its own device, but any stream may wait for any event.)
    CU(cudaSetDevice(src->device));
    CU(cudaStreamWaitEvent(src->h2d_stream, dst->build_ev, 0));
    return SVTME_OK;
}

extern "C" int svtme_release_picture(SvtMeHandle *h, uint64_t pic_num) {
    if (!h) return fail(SVTME_ERR_INVALID, "svtme_release_picture: null handle");
    std::lock_guard<std::mutex> lk(h->mu);
    auto it = h->index.find(pic_num);
    if (it == h->index.end()) return fail(SVTME_ERR_NOTFOUND, "svtme_release_picture: picture not resident");
    h->slots[it->second].used = false; // last_reader stays: the slot's next upload still waits for its readers
    h->index.erase(it);
    return SVTME_OK;
}

extern "C" int svtme_plane_device_ptr(SvtMeHandle *h, uint64_t pic_num, int level, void **ptr, size_t *bytes) {
    if (!h || level < 0 || level > 2 || !ptr) return fail(SVTME_ERR_INVALID, "svtme_plane_device_ptr: bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    const int si = find_slot(h, pic_num, false);
    if (si < 0) return fail(SVTME_ERR_NOTFOUND, "svtme_plane_device_ptr: picture not resident");
    *ptr = h->slots[si].pyr.lvl[level].base;
    if (bytes) *bytes = h->plane_bytes[level];
    return SVTME_OK;
}

extern "C" int svtme_get_plane(SvtMeHandle *h, uint64_t pic_num, int level, uint8_t *dst, uint32_t dst_stride, uint32_t pad) {
    if (!h || level < 0 || level > 2 || !dst) return fail(SVTME_ERR_INVALID, "svtme_get_plane: bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    CU(cudaSetDevice(h->device));
    const int si = find_slot(h, pic_num, false);
    if (si < 0) return fail(SVTME_ERR_NOTFOUND, "svtme_get_plane: picture not resident");
    const DevPlane &pl = h->slots[si].pyr.lvl[level];
    if ((int)pad > kPadT[level] || (int)pad > kPadL[level] || dst_stride < pl.width + 2 * pad)
        return fail(SVTME_ERR_INVALID, "svtme_get_plane: pad larger than the resident pad or stride too small");
    if (h->build_recorded) CU(cudaStreamWaitEvent(h->stream, h->build_ev, 0));
    CU(cudaMemcpy2DAsync(dst, dst_stride, pl.at(-(int)pad, -(int)pad), pl.stride, pl.width + 2 * pad, pl.height + 2 * pad,
                         cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return SVTME_OK;
}

// dg_detector_hme_level0 (pd_process.c:491) between two resident pictures; see svtme_b200.h
extern "C" int svtme_dg_detector_hme(SvtMeHandle *h, uint64_t src_pic, uint64_t ref_pic, uint32_t input_resolution, SvtMeDgMetrics *out) {
    if (!h || !out) return fail(SVTME_ERR_INVALID, "svtme_dg_detector_hme: null argument");
    if (input_resolution > 6) return fail(SVTME_ERR_INVALID, "svtme_dg_detector_hme: input_resolution is an EbInputResolution value (0..6)");
    std::lock_guard<std::mutex> lk(h->mu);
    CU(cudaSetDevice(h->device));
    const int ss = find_slot(h, src_pic, false), sr = find_slot(h, ref_pic, false);
    if (ss < 0 || sr < 0) return fail(SVTME_ERR_NOTFOUND, "svtme_dg_detector_hme: picture not resident");
    if (!h->dg_dev) {
        CU(cudaMalloc(&h->dg_dev, 256));
        const cudaError_t e = cudaMallocHost(&h->dg_host, 256);
        if (e != cudaSuccess) {
            cudaFree(h->dg_dev);
            h->dg_dev = nullptr;
            CU(e);
        }
    }
    if (h->build_recorded) CU(cudaStreamWaitEvent(h->stream, h->build_ev, 0));
    const int sa = input_resolution <= 1 ? 16 : (input_resolution <= 2 ? 64 : 128); // (:497-498)
    if (!h->stats_ev[0])
        for (cudaEvent_t &e : h->stats_ev) CU(cudaEventCreate(&e));
    CU(cudaEventRecord(h->stats_ev[0], h->stream));
    CU(svtme_launch_dg_hme(h->slots[ss].pyr.lvl[2], h->slots[sr].pyr.lvl[2], (int)h->b64_cols, (int)h->b64_rows, sa, h->dg_dev, h->stream));
    CU(cudaEventRecord(h->stats_ev[1], h->stream));
    CU(cudaMemcpyAsync(h->dg_host, h->dg_dev, svtme_dg_acc_bytes(), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaEventRecord(h->stats_ev[2], h->stream));
    CU(cudaStreamSynchronize(h->stream));
    svtme_dg_acc_unpack(h->dg_host, out);
    return SVTME_OK;
}

// svt_aom_gathering_picture_statistics (pic_analysis_process.c:1560) of resident pictures; see svtme_b200.h
extern "C" int svtme_pictures_statistics(SvtMeHandle *h, uint32_t n, const uint64_t *pic_nums, const SvtMeStatsParams *prm, SvtMeStatsOut *outs) {
    if (!h || !prm || (n && (!pic_nums || !outs))) return fail(SVTME_ERR_INVALID, "svtme_pictures_statistics: null argument");
    if (prm->calc_hist && (prm->regions_per_width < 1 || prm->regions_per_width > 4 || prm->regions_per_height < 1 || prm->regions_per_height > 4))
        return fail(SVTME_ERR_INVALID, "svtme_pictures_statistics: 1..4 regions per width / height (MAX_NUMBER_OF_REGIONS_IN_WIDTH / _HEIGHT)");
    for (uint32_t i = 0; i < n; ++i)
        if ((prm->calculate_variance && !outs[i].variance) || !outs[i].pic_avg_variance || !outs[i].avg_luma ||
            (prm->calc_hist && (!outs[i].picture_histogram || !outs[i].average_intensity_per_region)))
            return fail(SVTME_ERR_INVALID, "svtme_pictures_statistics: an output array the parameters ask for is null");
    if (!n) return SVTME_OK;
    std::lock_guard<std::mutex> lk(h->mu);
    CU(cudaSetDevice(h->device));
    std::vector<int> slot(n);
    for (uint32_t i = 0; i < n; ++i)
        if ((slot[i] = find_slot(h, pic_nums[i], false)) < 0) return fail(SVTME_ERR_NOTFOUND, "svtme_pictures_statistics: picture not resident");
    const int n_b64 = (int)(h->b64_cols * h->b64_rows);
    const size_t one = svtme_stats_scratch_bytes(n_b64), bytes = one * n;
    const size_t items_off = (bytes + 255) & ~(size_t)255, host_bytes = items_off + svtme_stats_item_bytes() * n;
    if (host_bytes > h->stats_cap) { // scratch and its pinned mirror (+ the item array) grow to the largest batch seen
        CU(cudaStreamSynchronize(h->stream));
        cudaFree(h->stats_dev);
        cudaFreeHost(h->stats_host);
        h->stats_dev = h->stats_host = nullptr;
        h->stats_cap = 0;
        CU(cudaMalloc(&h->stats_dev, host_bytes)); // (same layout on both sides: results, then the item array)
        const cudaError_t e = cudaMallocHost(&h->stats_host, host_bytes);
        if (e != cudaSuccess) {
            cudaFree(h->stats_dev);
            h->stats_dev = nullptr;
            CU(e);
        }
        h->stats_cap = host_bytes;
    }
    if (h->build_recorded) CU(cudaStreamWaitEvent(h->stream, h->build_ev, 0));
    size_t off_hist = 0, off_tail = 0;
    const size_t n_var = (size_t)n_b64 * (prm->all_block_sizes ? 85 : 1) * sizeof(uint16_t);
    for (uint32_t i = 0; i < n; ++i) svtme_stats_set_item(h->stats_host + items_off, (int)i, h->slots[slot[i]].pyr);
    if (!h->stats_ev[0])
        for (cudaEvent_t &e : h->stats_ev) CU(cudaEventCreate(&e));
    CU(cudaEventRecord(h->stats_ev[0], h->stream));
    // the item array goes to the device with one small copy (read through the mapped host pointer it cost every CTA a PCIe round trip)
    CU(cudaMemcpyAsync(h->stats_dev + items_off, h->stats_host + items_off, svtme_stats_item_bytes() * n, cudaMemcpyHostToDevice, h->stream));
    g_svtme_launches += svtme_launch_picture_statistics(h->slots[slot[0]].pyr, h->stats_dev + items_off, (int)n, *prm, h->stats_dev, &off_hist,
                                                        &off_tail, h->stream);
    CU(cudaGetLastError());
    CU(cudaEventRecord(h->stats_ev[1], h->stream));
    if (prm->calculate_variance)
        CU(cudaMemcpyAsync(h->stats_host, h->stats_dev, bytes, cudaMemcpyDeviceToHost, h->stream)); // rows, bins and tails of all n in one copy
    else
        for (uint32_t i = 0; i < n; ++i)
            CU(cudaMemcpyAsync(h->stats_host + one * i + off_hist, h->stats_dev + one * i + off_hist, one - off_hist, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaEventRecord(h->stats_ev[2], h->stream));
    CU(cudaStreamSynchronize(h->stream));
    for (uint32_t i = 0; i < n; ++i) {
        const uint8_t *src = h->stats_host + one * i;
        SvtMeStatsOut &o = outs[i];
        if (prm->calculate_variance) memcpy(o.variance, src, n_var);
        if (prm->calc_hist) memcpy(o.picture_histogram, src + off_hist, 16 * 256 * sizeof(uint32_t));
        unsigned long long avg_int[16], avg_luma = 0;
        svtme_stats_tail_unpack(src + off_tail, avg_int, &avg_luma, o.pic_avg_variance);
        if (prm->calc_hist)
            for (int k = 0; k < 16; ++k) o.average_intensity_per_region[k] = avg_int[k];
        *o.avg_luma = avg_luma;
    }
    return SVTME_OK;
}
extern "C" int svtme_picture_statistics(SvtMeHandle *h, uint64_t pic_num, const SvtMeStatsParams *prm, SvtMeStatsOut *out) {
    if (!out) return fail(SVTME_ERR_INVALID, "svtme_picture_statistics: null argument");
    return svtme_pictures_statistics(h, 1, &pic_num, prm, out);
}
extern "C" int svtme_last_statistics_ms(SvtMeHandle *h, float *kernels_ms, float *download_ms) {
    if (!h || !kernels_ms || !download_ms) return fail(SVTME_ERR_INVALID, "svtme_last_statistics_ms: null argument");
    std::lock_guard<std::mutex> lk(h->mu);
    if (!h->stats_ev[0]) return fail(SVTME_ERR_NOTFOUND, "svtme_last_statistics_ms: no statistics call yet");
    CU(cudaSetDevice(h->device));
    CU(cudaEventElapsedTime(kernels_ms, h->stats_ev[0], h->stats_ev[1]));
    CU(cudaEventElapsedTime(download_ms, h->stats_ev[1], h->stats_ev[2]));
    return SVTME_OK;
}
extern "C" uint32_t svtme_sizeof_stats_params(void) { return (uint32_t)sizeof(SvtMeStatsParams); }

extern "C" int svtme_sync(SvtMeHandle *h) {
    if (!h) return fail(SVTME_ERR_INVALID, "svtme_sync: null handle");
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->h2d_stream));
    CU(cudaStreamSynchronize(h->stream));
    for (int k = 1; k < SVTME_NS; ++k) CU(cudaStreamSynchronize(h->xstream[k]));
    CU(cudaStreamSynchronize(h->d2h_stream));
    return SVTME_OK;
}

// Upper bound of the full-pel search area of a picture (integer_search_b64 :1298-1320,1416-1441):
// min(sa_min * dist, sa_max), optional mv-based multiplier, width rounded up to 8, height >= 3,
// optional x1.5 of the 8x8-variance probe. Divisors only shrink it.
static bool areas_are_small(const SvtMeParams &p) {
    uint32_t w = p.me_sa.sa_max.width, hgt = p.me_sa.sa_max.height;
    if (p.mv_sa_adj_enabled) {
        const uint32_t mult = p.mv_sa_adj_sa_multiplier ? p.mv_sa_adj_sa_multiplier : 1;
        w *= mult;
        hgt *= mult;
    }
    w = ((w ? w : 1u) + 7u) & ~7u;
    hgt = hgt < 3 ? 3 : hgt;
    if (p.me_8x8_var_enabled && p.me_sr_mult2_th != 0xFFFFFFFFu) {
        w = ((w * 3 / 2) + 7u) & ~7u;
        hgt = hgt * 3 / 2;
    }
    return w * hgt <= (uint32_t)svtme_fp_direct_max();
}

// ---- result buffers: per job one packed record, same order on device and in the pinned mirror
struct OutLayout {
    size_t total, mv, cand, d64, d32, d16, d8, var8, rc, stat, gm, work, bsad, bmv, sc, hsad, doref, end;
};
static OutLayout out_layout(uint32_t n_b64, uint32_t n_pu, const SvtMeParams &p, bool deep) {
    OutLayout L;
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t r = o; o += (bytes + 15) & ~(size_t)15; return r; };
    L.total = take((size_t)n_b64 * n_pu);
    L.mv = take((size_t)n_b64 * n_pu * p.max_refs * 4);
    L.cand = take((size_t)n_b64 * n_pu * p.max_cand);
    L.d64 = take((size_t)n_b64 * 4); L.d32 = take((size_t)n_b64 * 4); L.d16 = take((size_t)n_b64 * 4);
    L.d8 = take((size_t)n_b64 * 4); L.var8 = take((size_t)n_b64 * 4); L.rc = take((size_t)n_b64 * 4);
    L.stat = take(n_b64); L.gm = take(n_b64);
    L.work = take(8 * SVTME_PHASES); // per launch slot
    if (deep) {
        L.bsad = take((size_t)n_b64 * 2 * 4 * 85 * 4); L.bmv = take((size_t)n_b64 * 2 * 4 * 85 * 4);
        L.sc = take((size_t)n_b64 * 2 * 4 * 2 * 2); L.hsad = take((size_t)n_b64 * 2 * 4 * 8); L.doref = take((size_t)n_b64 * 2 * 4);
    } else
        L.bsad = L.bmv = L.sc = L.hsad = L.doref = o;
    L.end = o;
    return L;
}
static bool wants_deep(const SvtMeOut &o) { return o.p_sb_best_sad || o.p_sb_best_mv || o.hme_sc || o.hme_sad || o.do_ref; }

static int validate_params(const SvtMeParams &p) {
    if (p.num_of_list_to_search < 1 || p.num_of_list_to_search > 2 || p.num_of_ref_pic_to_search[0] < 1 ||
        p.num_of_ref_pic_to_search[0] > 4 || p.num_of_ref_pic_to_search[1] > 4 || p.max_number_of_pus_per_sb != 85 ||
        p.num_hme_sa_w != 2 || p.num_hme_sa_h != 2 || p.max_refs == 0 || p.max_cand == 0)
        return fail(SVTME_ERR_INVALID, "svtme: unsupported SvtMeParams (lists/refs/PU count/HME regions)");
    if (p.num_of_list_to_search == 1 && p.num_of_ref_pic_to_search[1] != 0)
        return fail(SVTME_ERR_INVALID, "svtme: num_of_ref_pic_to_search[1] must be 0 for a single-list picture");
    return SVTME_OK;
}

// Enqueue the stage pipeline of n pictures on the main stream, results into batch B's device buffer.
static int launch_into(SvtMeHandle *h, Batch *B, const SvtMeJob *jobs, uint32_t n, int si = 0) {
    cudaStream_t st = si ? h->xstream[si] : h->stream;
    void *&d_states = si ? h->x_states[si] : h->d_states;
    size_t &states_bytes = si ? h->x_states_bytes[si] : h->states_bytes;
    if (B->jobs_pending) CU(cudaEventSynchronize(B->jobs_ev)); // the previous job array has left the pinned buffer
    B->jobs_pending = false;
    if (n > h->max_jobs) h->max_jobs = n;
    if (n > B->jobs_cap) {
        cudaFree(B->d_jobs);
        cudaFreeHost(B->h_jobs);
        B->d_jobs = B->h_jobs = nullptr;
        B->jobs_cap = 0;
        CU(cudaMalloc(&B->d_jobs, up16(sizeof(DevJob) * h->max_jobs)));
        CU(cudaMallocHost(&B->h_jobs, up16(sizeof(DevJob) * h->max_jobs)));
        B->jobs_cap = h->max_jobs;
    }
    size_t total = 0;
    for (uint32_t i = 0; i < n; ++i) {
        int rc = validate_params(jobs[i].params);
        if (rc) return rc;
        total += out_layout(h->n_b64, svtme_exported_pu_count(&jobs[i].params), jobs[i].params, wants_deep(jobs[i].out)).end;
    }
    if (total > B->ob.bytes) {
        // a queued batch's previous download has been collected before the batch is reused; the
        // synchronous batch may still be downloading only if the caller never fetched: wait for it
        CU(cudaStreamSynchronize(h->d2h_stream));
        CU(cudaStreamSynchronize(h->stream));
        for (int k = 1; k < SVTME_NS; ++k) CU(cudaStreamSynchronize(h->xstream[k]));
        cudaFree(B->ob.dev);
        cudaFreeHost(B->ob.host);
        B->ob = OutBuffers();
        if (total > h->max_ob_bytes) h->max_ob_bytes = total;
        CU(cudaMalloc(&B->ob.dev, h->max_ob_bytes));
        CU(cudaMallocHost(&B->ob.host, h->max_ob_bytes));
        B->ob.bytes = h->max_ob_bytes;
    }
    // every picture of the batch must be resident: checked before the sequence number and the slots' reader marks move
    // (a NOTFOUND half-way used to leave last_reader pointing at a launch event that is never recorded)
    for (uint32_t i = 0; i < n; ++i) {
        if (find_slot(h, jobs[i].cur, false) < 0) return fail(SVTME_ERR_NOTFOUND, "svtme_launch_pictures: current picture not resident");
        for (int li = 0; li < jobs[i].params.num_of_list_to_search; ++li)
            for (int ri = 0; ri < jobs[i].params.num_of_ref_pic_to_search[li]; ++ri)
                if (find_slot(h, jobs[i].refs[li][ri], false) < 0)
                    return fail(SVTME_ERR_NOTFOUND, "svtme_launch_pictures: reference picture not resident");
    }
    // sequence numbers name their stream: seq % SVTME_NS == si (the event ring is indexed by seq)
    do ++h->seq; while ((int)(h->seq % SVTME_NS) != si);
    B->seq = h->seq;
    B->n = n;
    B->total = total;
    size_t off = 0;
    B->work_off.clear();
    for (uint32_t i = 0; i < n; ++i) {
        const SvtMeJob &j = jobs[i];
        DevJob &d = B->h_jobs[i];
        memset(&d, 0, sizeof(d));
        d.prm = j.params;
        const int sc = find_slot(h, j.cur, false);
        if (sc < 0) return fail(SVTME_ERR_NOTFOUND, "svtme_launch_pictures: current picture not resident");
        d.cur = h->slots[sc].pyr;
        h->slots[sc].last_reader[si] = B->seq;
        for (int li = 0; li < j.params.num_of_list_to_search; ++li)
            for (int ri = 0; ri < j.params.num_of_ref_pic_to_search[li]; ++ri) {
                const int sr = find_slot(h, j.refs[li][ri], false);
                if (sr < 0) return fail(SVTME_ERR_NOTFOUND, "svtme_launch_pictures: reference picture not resident");
                d.ref[li][ri] = h->slots[sr].pyr;
                h->slots[sr].last_reader[si] = B->seq;
            }
        d.width = (int)h->width;
        d.height = (int)h->height;
        d.b64_cols = (int)h->b64_cols;
        d.b64_rows = (int)h->b64_rows;
        d.n_pu = (int)svtme_exported_pu_count(&j.params);
        const bool deep = wants_deep(j.out);
        const OutLayout L = out_layout(h->n_b64, d.n_pu, j.params, deep);
        uint8_t *b = B->ob.dev + off;
        d.total = b + L.total;
        d.mv = (uint32_t *)(b + L.mv);
        d.cand = b + L.cand;
        d.dist64 = (uint32_t *)(b + L.d64); d.dist32 = (uint32_t *)(b + L.d32); d.dist16 = (uint32_t *)(b + L.d16);
        d.dist8 = (uint32_t *)(b + L.d8); d.var8 = (uint32_t *)(b + L.var8); d.rc_dist = (uint32_t *)(b + L.rc);
        d.stationary = b + L.stat; d.allow_gm = b + L.gm;
        d.work_ops = (unsigned long long *)(b + L.work);
        if (deep) {
            d.best_sad = (uint32_t *)(b + L.bsad); d.best_mv = (uint32_t *)(b + L.bmv);
            d.hme_sc = (int16_t *)(b + L.sc); d.hme_sad = (unsigned long long *)(b + L.hsad); d.do_ref = b + L.doref;
        }
        B->work_off.push_back(off + L.work);
        off += L.end;
    }
    // entries the stage never writes (pruned candidates) read as zero, like a fresh MeSbResults
    if (h->build_recorded) CU(cudaStreamWaitEvent(st, h->build_ev, 0)); // pyramids are built on the upload stream
    if (h->ext_recorded) CU(cudaStreamWaitEvent(st, h->ext_ev, 0));     // ... or written by an exchange on the caller's stream
    CU(svtme_launch_zero_fill(B->ob.dev, total, st));
    CU(svtme_launch_fetch_from_host(B->d_jobs, B->h_jobs, up16(sizeof(DevJob) * n), st));
    CU(cudaEventRecord(B->jobs_ev, st));
    B->jobs_pending = true;
    CU(cudaEventRecord(B->ev0, st));
    {
        const size_t need = up16((size_t)n * h->n_b64 * svtme_stage_state_bytes());
        if (need > states_bytes) {
            CU(cudaStreamSynchronize(st)); // earlier launches still use the old scratch
            cudaFree(d_states);
            d_states = nullptr;
            states_bytes = 0;
            CU(cudaMalloc(&d_states, need));
            states_bytes = need;
        }
    }
    SvtmeStageFlags flags = {};
    flags.small_areas = true;
    flags.prehme_strip = true;
    flags.refine_small = getenv("SVTME_NO_REFINE_KERNEL") == nullptr; // (tests exercise both paths)
    flags.l0_strip = getenv("SVTME_NO_L0_STRIP") == nullptr;
    for (uint32_t i = 0; i < n; ++i) {
        flags.prehme_strip = flags.prehme_strip && jobs[i].params.hme_search_method != 1;
        const SvtmeStageFlags f = svtme_stage_flags(jobs[i].params);
        flags.small_areas = flags.small_areas && areas_are_small(jobs[i].params);
        flags.refine_small = flags.refine_small && f.refine_small;
        flags.int_dependent |= f.int_dependent;
        for (int li = 0; li < 2; ++li) flags.max_refs[li] = f.max_refs[li] > flags.max_refs[li] ? f.max_refs[li] : flags.max_refs[li];
        flags.zz |= f.zz; flags.prehme |= f.prehme; flags.prehme_l1_round |= f.prehme_l1_round;
        flags.hme_l0 |= f.hme_l0; flags.hme_l0_round_b |= f.hme_l0_round_b;
        flags.hme_l1 |= f.hme_l1; flags.hme_l2 |= f.hme_l2;
    }
    B->sev_mask = 0;
    if (h->search_timing && !B->sev[0])
        for (int k = 0; k < 4 * SVTME_PHASES; ++k) CU(cudaEventCreate(&B->sev[k]));
    svtme_launch_me_stage(B->d_jobs, (int)n, (int)h->b64_cols, (int)h->b64_rows, d_states, st, B->ph_ev, flags,
                          h->search_timing ? B->sev : nullptr, &B->sev_mask);
    CU(cudaGetLastError());
    CU(cudaEventRecord(B->ev1, st));
    CU(cudaEventRecord(h->launch_ev[B->seq % SVTME_LAUNCH_EV_RING], st));
    h->last = B;
    return SVTME_OK;
}

// Scatter batch B's pinned result mirror into the callers' SvtMeOut arrays.
static int scatter_results(SvtMeHandle *h, const Batch *B, const SvtMeJob *jobs, uint32_t n) {
    std::vector<size_t> offs(n);
    size_t off = 0;
    for (uint32_t i = 0; i < n; ++i) {
        offs[i] = off;
        off += out_layout(h->n_b64, svtme_exported_pu_count(&jobs[i].params), jobs[i].params, wants_deep(jobs[i].out)).end;
    }
    // plain memcpy of ~0.4 MB per picture; a few host threads when the batch is large (one thread moves ~10 GB/s).
    // SVTME_SCATTER_THREADS (1..8, default 4) lets a host that runs one process per GPU keep its cores.
    static const int scatter_threads = [] {
        const char *e = getenv("SVTME_SCATTER_THREADS");
        const int v = e ? atoi(e) : 4;
        return v < 1 ? 1 : (v > 8 ? 8 : v);
    }();
#pragma omp parallel for schedule(static) num_threads(scatter_threads) if (n >= 4 && scatter_threads > 1)
    for (int i = 0; i < (int)n; ++i) {
        const SvtMeJob &j = jobs[i];
        const uint32_t n_pu = svtme_exported_pu_count(&j.params), nb = h->n_b64;
        const bool deep = wants_deep(j.out);
        const OutLayout L = out_layout(nb, n_pu, j.params, deep);
        const uint8_t *b = B->ob.host + offs[i];
#define CP(dst, o, bytes) if (j.out.dst) memcpy(j.out.dst, b + (o), (bytes))
        if (!(i < (int)B->direct.size() && B->direct[i])) { // (else: already in place, see svtme_submit_pictures)
            CP(total_me_candidate_index, L.total, (size_t)nb * n_pu);
            CP(me_mv_array, L.mv, (size_t)nb * n_pu * j.params.max_refs * 4);
            CP(me_candidate_array, L.cand, (size_t)nb * n_pu * j.params.max_cand);
        }
        CP(me_64x64_distortion, L.d64, (size_t)nb * 4); CP(me_32x32_distortion, L.d32, (size_t)nb * 4);
        CP(me_16x16_distortion, L.d16, (size_t)nb * 4); CP(me_8x8_distortion, L.d8, (size_t)nb * 4);
        CP(me_8x8_cost_variance, L.var8, (size_t)nb * 4); CP(rc_me_distortion, L.rc, (size_t)nb * 4);
        CP(stationary_block_present_sb, L.stat, nb); CP(rc_me_allow_gm, L.gm, nb);
        if (deep) {
            CP(p_sb_best_sad, L.bsad, (size_t)nb * 2 * 4 * 85 * 4); CP(p_sb_best_mv, L.bmv, (size_t)nb * 2 * 4 * 85 * 4);
            CP(hme_sc, L.sc, (size_t)nb * 2 * 4 * 2 * 2); CP(hme_sad, L.hsad, (size_t)nb * 2 * 4 * 8); CP(do_ref, L.doref, (size_t)nb * 2 * 4);
        }
#undef CP
    }
    return SVTME_OK;
}

static size_t results_bytes(const SvtMeHandle *h, const SvtMeJob *jobs, uint32_t n) {
    size_t total = 0;
    for (uint32_t i = 0; i < n; ++i)
        total += out_layout(h->n_b64, svtme_exported_pu_count(&jobs[i].params), jobs[i].params, wants_deep(jobs[i].out)).end;
    return total;
}

extern "C" int svtme_launch_pictures(SvtMeHandle *h, const SvtMeJob *jobs, uint32_t n) {
    if (!h || !jobs || n == 0) return fail(SVTME_ERR_INVALID, "svtme_launch_pictures: bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    CU(cudaSetDevice(h->device));
    return launch_into(h, h->sync_batch, jobs, n);
}

extern "C" int svtme_fetch_results(SvtMeHandle *h, const SvtMeJob *jobs, uint32_t n) {
    if (!h || !jobs || n == 0) return fail(SVTME_ERR_INVALID, "svtme_fetch_results: bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    CU(cudaSetDevice(h->device));
    Batch *B = h->sync_batch;
    if (B->n != n || results_bytes(h, jobs, n) != B->total)
        return fail(SVTME_ERR_INVALID, "svtme_fetch_results: does not match the last svtme_launch_pictures");
    CU(cudaMemcpyAsync(B->ob.host, B->ob.dev, B->total, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    cudaEventElapsedTime(&h->last_ms, B->ev0, B->ev1);
    return scatter_results(h, B, jobs, n);
}

extern "C" int svtme_run_pictures(SvtMeHandle *h, const SvtMeJob *jobs, uint32_t n) {
    int rc = svtme_launch_pictures(h, jobs, n);
    if (rc) return rc;
    return svtme_fetch_results(h, jobs, n);
}

// ---- queued form: several batches in flight, results collected in submission order. The download of
// a batch runs on its own stream as soon as the batch's pipeline is done, i.e. while the next batch
// computes and the one after that uploads (svtme_put_pictures uses a third stream).
extern "C" int svtme_submit_pictures(SvtMeHandle *h, const SvtMeJob *jobs, uint32_t n) {
    if (!h || !jobs || n == 0) return fail(SVTME_ERR_INVALID, "svtme_submit_pictures: bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    CU(cudaSetDevice(h->device));
    // a recycled batch object whose buffers already hold this batch (best fit), else the largest free one (it grows),
    // else a new one: growing a buffer synchronises the device, which a steady stream of batches must not do
    Batch *B = nullptr;
    if (!h->free_batches.empty()) {
        const size_t need = results_bytes(h, jobs, n);
        size_t best = 0;
        bool fits = false;
        for (size_t i = 0; i < h->free_batches.size(); ++i) {
            const Batch *c = h->free_batches[i];
            const bool ok = c->ob.bytes >= need && c->jobs_cap >= n;
            const Batch *b = h->free_batches[best];
            if (ok ? (!fits || c->ob.bytes < b->ob.bytes) : (!fits && c->ob.bytes > b->ob.bytes)) best = i;
            fits |= ok;
        }
        B = h->free_batches[best];
        h->free_batches.erase(h->free_batches.begin() + (long)best);
    } else if (!(B = batch_new()))
        return fail(SVTME_ERR_NOMEM, "svtme_submit_pictures: batch allocation");
    int rc = launch_into(h, B, jobs, n, (int)(h->submit_count++ % (uint32_t)h->n_streams));
    if (rc) {
        h->free_batches.push_back(B);
        return rc;
    }
    CU(cudaStreamWaitEvent(h->d2h_stream, B->ev1, 0));
    // Download. The three big tables of a job (candidate counts, motion vectors, candidates: 99 % of the bytes) go straight
    // into the caller's arrays when those are page-locked (cudaMallocHost / svtme_pin_host_memory): no staging copy, no
    // host-side scatter — at 4K preset 4 the scatter of 1.08 GB per 120-frame step was as long as the ME itself. Everything
    // else (and every job whose arrays are pageable) travels through the batch's pinned mirror as before.
    B->direct.assign(n, 0);
    {
        size_t off = 0, run0 = 0; // [run0, off): bytes not yet copied to the mirror
        for (uint32_t i = 0; i < n; ++i) {
            const SvtMeJob &j = jobs[i];
            const uint32_t n_pu = svtme_exported_pu_count(&j.params), nb = h->n_b64;
            const OutLayout L = out_layout(nb, n_pu, j.params, wants_deep(j.out));
            bool pinned = j.out.total_me_candidate_index && j.out.me_mv_array && j.out.me_candidate_array;
            const void *big[3] = {j.out.total_me_candidate_index, j.out.me_mv_array, j.out.me_candidate_array};
            for (int k = 0; k < 3 && pinned; ++k) {
                cudaPointerAttributes at;
                if (cudaPointerGetAttributes(&at, big[k]) != cudaSuccess || at.type != cudaMemoryTypeHost) pinned = false;
            }
            (void)cudaGetLastError(); // (an unregistered host pointer makes older runtimes report an error: not one of ours)
            if (pinned) {
                if (off + L.total > run0)
                    CU(cudaMemcpyAsync(B->ob.host + run0, B->ob.dev + run0, off + L.total - run0, cudaMemcpyDeviceToHost, h->d2h_stream));
                const uint8_t *d = B->ob.dev + off;
                CU(cudaMemcpyAsync(j.out.total_me_candidate_index, d + L.total, (size_t)nb * n_pu, cudaMemcpyDeviceToHost, h->d2h_stream));
                CU(cudaMemcpyAsync(j.out.me_mv_array, d + L.mv, (size_t)nb * n_pu * j.params.max_refs * 4, cudaMemcpyDeviceToHost, h->d2h_stream));
                CU(cudaMemcpyAsync(j.out.me_candidate_array, d + L.cand, (size_t)nb * n_pu * j.params.max_cand, cudaMemcpyDeviceToHost,
                                   h->d2h_stream));
                B->direct[i] = 1;
                run0 = off + L.d64; // the mirror run resumes behind the three tables
            }
            off += L.end;
        }
        if (off > run0) CU(cudaMemcpyAsync(B->ob.host + run0, B->ob.dev + run0, off - run0, cudaMemcpyDeviceToHost, h->d2h_stream));
    }
    CU(cudaEventRecord(B->done_ev, h->d2h_stream));
    h->queue.push_back(B);
    return SVTME_OK;
}

extern "C" int svtme_collect_results(SvtMeHandle *h, const SvtMeJob *jobs, uint32_t n) {
    if (!h || !jobs || n == 0) return fail(SVTME_ERR_INVALID, "svtme_collect_results: bad argument");
    Batch *B;
    {
        std::lock_guard<std::mutex> lk(h->mu);
        CU(cudaSetDevice(h->device));
        if (h->queue.empty()) return fail(SVTME_ERR_NOTFOUND, "svtme_collect_results: nothing submitted");
        B = h->queue.front();
        if (B->n != n || results_bytes(h, jobs, n) != B->total)
            return fail(SVTME_ERR_INVALID, "svtme_collect_results: jobs do not match the oldest submitted batch");
        h->queue.pop_front();
    }
    // wait and scatter outside the lock: other threads keep submitting
    cudaError_t e = cudaEventSynchronize(B->done_ev);
    int rc = e == cudaSuccess ? scatter_results(h, B, jobs, n) : fail(SVTME_ERR_CUDA, "svtme_collect_results", e);
    std::lock_guard<std::mutex> lk(h->mu);
    if (e == cudaSuccess) cudaEventElapsedTime(&h->last_ms, B->ev0, B->ev1);
    if (e == cudaSuccess && h->origin_ev) {
        cudaEventElapsedTime(&h->span[0], h->origin_ev, B->ev0);
        cudaEventElapsedTime(&h->span[1], h->origin_ev, B->ev1);
        cudaEventElapsedTime(&h->span[2], h->origin_ev, B->done_ev);
    }
    h->free_batches.push_back(B);
    return rc;
}

// Timeline aid: svtme_debug_mark records an origin on the main stream; after each svtme_collect_results,
// svtme_debug_last_span returns when that batch's pipeline started / ended and when its download ended (ms after the mark).
extern "C" int svtme_debug_mark(SvtMeHandle *h) {
    if (!h) return fail(SVTME_ERR_INVALID, "svtme_debug_mark: null handle");
    std::lock_guard<std::mutex> lk(h->mu);
    CU(cudaSetDevice(h->device));
    if (!h->origin_ev) CU(cudaEventCreate(&h->origin_ev));
    CU(cudaEventRecord(h->origin_ev, h->stream));
    return SVTME_OK;
}
extern "C" int svtme_debug_last_span(SvtMeHandle *h, float *ms3) {
    if (!h || !ms3) return fail(SVTME_ERR_INVALID, "svtme_debug_last_span: bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    for (int k = 0; k < 3; ++k) ms3[k] = h->span[k];
    return SVTME_OK;
}

// Page-lock caller memory in place (cudaHostRegister) so that results can be downloaded straight into it.
extern "C" int svtme_pin_host_memory(void *ptr, size_t bytes) {
    if (!ptr || !bytes) return fail(SVTME_ERR_INVALID, "svtme_pin_host_memory: bad argument");
    CU(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
    return SVTME_OK;
}
extern "C" int svtme_unpin_host_memory(void *ptr) {
    if (!ptr) return fail(SVTME_ERR_INVALID, "svtme_unpin_host_memory: bad argument");
    CU(cudaHostUnregister(ptr));
    return SVTME_OK;
}

extern "C" uint32_t svtme_pending_batches(SvtMeHandle *h) {
    if (!h) return 0;
    std::lock_guard<std::mutex> lk(h->mu);
    return (uint32_t)h->queue.size();
}

extern "C" int svtme_run_picture(SvtMeHandle *h, const SvtMeParams *params, uint64_t cur,
                                 const uint64_t refs[SVTME_MAX_LISTS][SVTME_MAX_REFS], SvtMeOut *out) {
    if (!params || !refs || !out) return fail(SVTME_ERR_INVALID, "svtme_run_picture: null argument");
    SvtMeJob j;
    memset(&j, 0, sizeof(j));
    j.params = *params;
    j.cur = cur;
    memcpy(j.refs, refs, sizeof(j.refs));
    j.out = *out;
    return svtme_run_pictures(h, &j, 1);
}

extern "C" float svtme_last_kernel_ms(SvtMeHandle *h) { return h ? h->last_ms : 0.f; }

// Run the handle's work on a caller-owned CUDA stream (e.g. torch's current stream, so that
// torch.cuda.Event timing brackets these kernels). Pass NULL to return to the private stream.
extern "C" int svtme_set_stream(SvtMeHandle *h, void *cuda_stream) {
    if (!h) return fail(SVTME_ERR_INVALID, "svtme_set_stream: null handle");
    std::lock_guard<std::mutex> lk(h->mu);
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    if (h->own_stream) cudaStreamDestroy(h->stream);
    if (cuda_stream) {
        h->stream = (cudaStream_t)cuda_stream;
        h->own_stream = false;
    } else {
        CU(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
        h->own_stream = true;
    }
    return SVTME_OK;
}

// Sum of the algorithmic SAD byte-ops (positions x block pixels actually compared) the stage
// kernel executed in the last svtme_launch_pictures call. Synchronises the stream.
extern "C" int svtme_last_work_ops(SvtMeHandle *h, uint64_t *ops) {
    if (!h || !ops) return fail(SVTME_ERR_INVALID, "svtme_last_work_ops: bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    for (int k = 1; k < SVTME_NS; ++k) CU(cudaStreamSynchronize(h->xstream[k]));
    uint64_t sum = 0;
    if (!h->last) return fail(SVTME_ERR_NOTFOUND, "svtme_last_work_ops: nothing launched");
    for (size_t o : h->last->work_off) {
        unsigned long long v[SVTME_PHASES];
        CU(cudaMemcpy(v, h->last->ob.dev + o, sizeof(v), cudaMemcpyDeviceToHost));
        for (int k = 0; k < SVTME_PHASES; ++k) sum += v[k];
    }
    *ops = sum;
    cudaEventElapsedTime(&h->last_ms, h->last->ev0, h->last->ev1);
    return SVTME_OK;
}

// Search launches of each slot timed on their own (see svtme_b200.h).
extern "C" int svtme_set_search_timing(SvtMeHandle *h, int on) {
    if (!h) return fail(SVTME_ERR_INVALID, "svtme_set_search_timing: null handle");
    std::lock_guard<std::mutex> lk(h->mu);
    h->search_timing = on != 0;
    return SVTME_OK;
}

extern "C" int svtme_last_search_ms(SvtMeHandle *h, float *ms, uint32_t n) {
    if (!h || !ms) return fail(SVTME_ERR_INVALID, "svtme_last_search_ms: bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    for (int k = 1; k < SVTME_NS; ++k) CU(cudaStreamSynchronize(h->xstream[k]));
    if (n > SVTME_PHASES) n = SVTME_PHASES;
    if (!h->last) return fail(SVTME_ERR_NOTFOUND, "svtme_last_search_ms: nothing launched");
    for (uint32_t k = 0; k < n; ++k) {
        ms[k] = 0.f;
        for (int r = 0; r < 2; ++r)
            if (h->last->sev_mask >> (2 * k + r) & 1) {
                float t = 0.f;
                CU(cudaEventElapsedTime(&t, h->last->sev[(2 * k + r) * 2], h->last->sev[(2 * k + r) * 2 + 1]));
                ms[k] += t;
            }
    }
    return SVTME_OK;
}

// Algorithmic SAD byte-ops per launch slot (same slots as svtme_last_phase_ms) of the last launch.
extern "C" int svtme_last_phase_work(SvtMeHandle *h, uint64_t *ops, uint32_t n) {
    if (!h || !ops) return fail(SVTME_ERR_INVALID, "svtme_last_phase_work: bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    for (int k = 1; k < SVTME_NS; ++k) CU(cudaStreamSynchronize(h->xstream[k]));
    if (n > SVTME_PHASES) n = SVTME_PHASES;
    for (uint32_t k = 0; k < n; ++k) ops[k] = 0;
    if (!h->last) return fail(SVTME_ERR_NOTFOUND, "svtme_last_phase_work: nothing launched");
    for (size_t o : h->last->work_off) {
        unsigned long long v[SVTME_PHASES];
        CU(cudaMemcpy(v, h->last->ob.dev + o, sizeof(v), cudaMemcpyDeviceToHost));
        for (uint32_t k = 0; k < n; ++k) ops[k] += v[k];
    }
    return SVTME_OK;
}

// Device time of each launch slot of the stage pipeline in the last svtme_launch_pictures call
// (names: SVTME_PHASE_NAMES). Synchronises the stream.
extern "C" int svtme_last_phase_ms(SvtMeHandle *h, float *ms, uint32_t n) {
    if (!h || !ms) return fail(SVTME_ERR_INVALID, "svtme_last_phase_ms: bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    for (int k = 1; k < SVTME_NS; ++k) CU(cudaStreamSynchronize(h->xstream[k]));
    if (n > SVTME_PHASES) n = SVTME_PHASES;
    if (!h->last) return fail(SVTME_ERR_NOTFOUND, "svtme_last_phase_ms: nothing launched");
    for (uint32_t k = 0; k < n; ++k) {
        ms[k] = 0.f;
        CU(cudaEventElapsedTime(&ms[k], k ? h->last->ph_ev[k - 1] : h->last->ev0, h->last->ph_ev[k]));
    }
    return SVTME_OK;
}
