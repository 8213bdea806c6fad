// svtme_noalloc.cu — svt_sad_loop_kernel_b200_noalloc: a batched svt_sad_loop_kernel for callers that must not
// allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// NO COUNTERPART IN THE REFERENCE. It is a variant of the kernel-test
// wrapper svt_sad_loop_kernel_b200 (svtme_testabi.cu), which cudaMallocs its inputs and an 8-byte result key per
// call. Here the caller passes DEVICE inputs and outputs and the per-search result keys — the only scratch memory
// the kernel needs — live in a file-scope `static __device__` array of SVTME_NOALLOC_MAX_SEARCHES entries (64 KB,
// held by every GPU that loads this library until the process exits).
//
//   * The entry point calls no cudaMalloc / cudaFree / allocating library routine; it only launches two kernels on
//     the caller's stream. Set-up (svtme_noalloc_setup: the scratch array's device address through
//     cudaGetSymbolAddress — in host code the array's NAME is a host-side stand-in, not a device address — and the
//     kernel's dynamic shared-memory attribute) runs once per device with that device current; call it ahead of
//     time where even lazy module loading is unwelcome, otherwise the first call on a device does it.
//   * Every call shares the one array, which keeps the last call's contents. Nothing has to start at zero: each
//     call first overwrites entries [0, n_searches) and only then reads them. Calls that overlap on the GPU (two
//     streams, two host threads on one device) RACE on the array: serialise them on one stream.
#include <stdio.h>
#include <mutex>
#include "svtme_common.cuh"
namespace { // this file's own copy of the search engine (header-only device code): internal linkage, so its
            // non-inline functions do not collide with svtme_stage.cu's copy at link time
#include "svtme_search.cuh"
} // namespace

void svtme_set_error(const char *what); // svtme_api.cu: text behind svtme_last_error()

static __device__ __align__(256) unsigned long long g_noalloc_keys[SVTME_NOALLOC_MAX_SEARCHES];

namespace {
struct NoallocSmem {
    HJob job;
    HPlan plan;
    __align__(16) uint8_t src[64 * 64];
    __align__(16) uint8_t arena[SVTME_HME_ARENA_BYTES];
};

// One CTA per search: source block -> shared memory, then the engine every svt_sad_loop_kernel call of the stage
// uses (hme_search_multi, svtme_search.cuh); the winner's key  sad << 32 | y << 16 | x  goes to keys[search].
__global__ void __launch_bounds__(SVTME_HME_THREADS) sad_loop_batch_kernel(const uint8_t *src, int src_stride, size_t src_step,
                                                                           const uint8_t *ref, int ref_stride, size_t ref_step,
                                                                           int bh, int bw, int raw, int skip, int sa_w,
                                                                           int sa_h, unsigned long long *keys) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    NoallocSmem *sm = reinterpret_cast<NoallocSmem *>(smem_raw);
    const uint8_t *s = src + (size_t)blockIdx.x * src_step;
    for (int i = threadIdx.x; i < 64 * 64; i += (int)blockDim.x) {
        const int r = i >> 6, c = i & 63;
        sm->src[i] = (r < bh && c < bw) ? s[(size_t)r * src_stride + c] : 0;
    }
    const int row_step = ref_stride / raw; // the RTCD caller multiplies both strides to sub-sample rows
    if (threadIdx.x == 0) {
        sm->job.ref = ref + (size_t)blockIdx.x * ref_step;
        sm->job.stride = raw;
        sm->job.sa_w = sa_w;
        sm->job.sa_h = sa_h;
        sm->job.skip_even = (bw == 16 && bh <= 16 && skip) ? 1 : 0;
        sm->job.key = HME_KEY_INIT;
    }
    __syncthreads();
    const HBlk blk = make_hblk(sm->src, 64, bw, bh, row_step);
    hme_search_multi(blk, &sm->job, 1, &sm->plan, sm->arena);
    if (threadIdx.x == 0) keys[blockIdx.x] = sm->job.key;
}

// keys -> the RTCD outputs: best_sad = 0xffffff and x / y untouched when no position beat the initial best
// (compute_sad_c.c:58-101).
__global__ void sad_loop_decode_kernel(const unsigned long long *keys, uint32_t n, unsigned long long *best_sad, int16_t *x,
                                       int16_t *y) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long key = keys[i];
    if ((key >> 32) < 0xffffffULL) {
        best_sad[i] = key >> 32;
        x[i] = (int16_t)(key & 0xffff);
        y[i] = (int16_t)((key >> 16) & 0xffff);
    } else
        best_sad[i] = 0xffffff;
}

std::mutex g_setup_mu;
unsigned long long *g_keys_dev[64] = {}; // device address of g_noalloc_keys, per device
} // namespace

static int noalloc_fail(const char *what, cudaError_t e = cudaSuccess) {
    char msg[384];
    if (e != cudaSuccess)
        snprintf(msg, sizeof(msg), "%s: %s", what, cudaGetErrorString(e));
    else
        snprintf(msg, sizeof(msg), "%s", what);
    svtme_set_error(msg);
    return e != cudaSuccess ? SVTME_ERR_CUDA : SVTME_ERR_INVALID;
}

extern "C" int svtme_noalloc_setup(int device) {
    if (device < 0 || device >= 64) return noalloc_fail("svtme_noalloc_setup: device index out of range");
    std::lock_guard<std::mutex> lk(g_setup_mu);
    if (g_keys_dev[device]) return SVTME_OK;
    int cur = -1;
    cudaError_t e = cudaGetDevice(&cur);
    if (e != cudaSuccess) return noalloc_fail("svtme_noalloc_setup: cudaGetDevice", e);
    if (cur != device && (e = cudaSetDevice(device)) != cudaSuccess) return noalloc_fail("svtme_noalloc_setup: cudaSetDevice", e);
    unsigned long long *ptr = nullptr;
    e = cudaGetSymbolAddress((void **)&ptr, g_noalloc_keys); // with `device` current: the address is per device
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(sad_loop_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(NoallocSmem));
    if (cur != device && cur >= 0) cudaSetDevice(cur);
    if (e != cudaSuccess || !ptr) return noalloc_fail("svtme_noalloc_setup: cudaGetSymbolAddress / cudaFuncSetAttribute", e);
    g_keys_dev[device] = ptr;
    return SVTME_OK;
}

extern "C" int svt_sad_loop_kernel_b200_noalloc(void *cuda_stream, uint32_t n_searches, const uint8_t *d_src, uint32_t src_stride,
                                                size_t src_step, const uint8_t *d_ref, uint32_t ref_stride, size_t ref_step,
                                                uint32_t block_height, uint32_t block_width, uint32_t src_stride_raw,
                                                uint8_t skip_search_line, int16_t search_area_width,
                                                int16_t search_area_height, uint64_t *d_best_sad, int16_t *d_x_search_center,
                                                int16_t *d_y_search_center) {
    if (n_searches == 0) return SVTME_OK;
    if (n_searches > SVTME_NOALLOC_MAX_SEARCHES) {
        char msg[160];
        snprintf(msg, sizeof(msg), "svt_sad_loop_kernel_b200_noalloc: %u searches exceed the static scratch array (maximum %d per call)",
                 n_searches, SVTME_NOALLOC_MAX_SEARCHES);
        return noalloc_fail(msg);
    }
    if (!d_src || !d_ref || !d_best_sad || !d_x_search_center || !d_y_search_center)
        return noalloc_fail("svt_sad_loop_kernel_b200_noalloc: null pointer");
    if (block_width == 0 || block_height == 0 || block_width > 64 || block_height > 64 || src_stride_raw == 0 ||
        ref_stride % src_stride_raw || search_area_width <= 0 || search_area_height <= 0)
        return noalloc_fail("svt_sad_loop_kernel_b200_noalloc: unsupported geometry (block 1..64 x 1..64, ref_stride = k * "
                            "src_stride_raw, search area >= 1x1)");
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) return noalloc_fail("svt_sad_loop_kernel_b200_noalloc: cudaGetDevice", e);
    if (!g_keys_dev[dev]) {
        const int rc = svtme_noalloc_setup(dev);
        if (rc) return rc;
    }
    unsigned long long *keys = g_keys_dev[dev]; // NOT g_noalloc_keys: that name is only valid in device code
    cudaStream_t st = (cudaStream_t)cuda_stream;
    sad_loop_batch_kernel<<<n_searches, SVTME_HME_THREADS, sizeof(NoallocSmem), st>>>(
        d_src, (int)src_stride, src_step, d_ref, (int)ref_stride, ref_step, (int)block_height, (int)block_width, (int)src_stride_raw,
        skip_search_line, search_area_width, search_area_height, keys);
    sad_loop_decode_kernel<<<(n_searches + 127) / 128, 128, 0, st>>>(keys, n_searches, reinterpret_cast<unsigned long long *>(d_best_sad),
                                                                     d_x_search_center, d_y_search_center);
    g_svtme_launches += 2;
    e = cudaGetLastError();
    if (e != cudaSuccess) return noalloc_fail("svt_sad_loop_kernel_b200_noalloc: launch", e);
    return SVTME_OK;
}
