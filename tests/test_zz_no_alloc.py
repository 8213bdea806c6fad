"""GPU test (-m gpu) of svt_sad_loop_kernel_b200_noalloc — the allocation-free batched svt_sad_loop_kernel whose
scratch memory is a static device array (include/svtme_b200.h; no counterpart in the reference).
Its outputs must equal, element for element and bit for bit (integer SADs: no
tolerance), what the existing entry point svt_sad_loop_kernel_b200 returns for the same search, on two calls in a
row (the second with other inputs and fewer searches: the array keeps the first call's contents), through the C
ABI. The file name makes pytest collect it last; it passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

import _util as U

pytestmark = pytest.mark.gpu
vp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
MAX_SEARCHES = 8192  # SVTME_NOALLOC_MAX_SEARCHES


def existing_entry_point(lib, src, src_stride, ref, ref_off, ref_stride, bh, bw, raw, skip, sw, sh):
    """svt_sad_loop_kernel_b200 (host buffers; allocates device memory itself) for one search."""
    best, x, y = C.c_uint64(0), C.c_int16(-3), C.c_int16(-5)
    lib.svt_sad_loop_kernel_b200(vp(src), C.c_uint32(src_stride), C.c_void_p(ref.ctypes.data + ref_off), C.c_uint32(ref_stride),
                                 C.c_uint32(bh), C.c_uint32(bw), C.byref(best), C.byref(x), C.byref(y), C.c_uint32(raw),
                                 C.c_uint8(skip), C.c_int16(sw), C.c_int16(sh))
    return best.value, x.value, y.value


def run_case(lib, torch, rng, n, bw, bh, sw, sh, sub, skip):
    """n searches of one geometry: block i is a tile of `blocks`, window i starts at column 16 + 5 * i of `ref`."""
    raw = 1024                                  # reference plane stride
    src_stride, src_step = 64 * sub, 64 * 64 * sub
    blocks = rng.integers(0, 256, size=(n, 64 * sub, 64), dtype=np.uint8)
    rows = 16 + sh + 64 * sub + 8               # front / tail rows: the kernel may read 16 B before and 4 rows past a window
    ref = rng.integers(0, 256, size=(rows, raw), dtype=np.uint8)
    ref_off0, ref_step = 8 * raw + 16, 5
    assert ref_off0 % raw + ref_step * (n - 1) + sw + bw + 256 < raw
    d_blocks, d_ref = torch.from_numpy(blocks).cuda(), torch.from_numpy(ref).cuda()
    d_best = torch.zeros(n, dtype=torch.int64, device="cuda")
    d_x = torch.full((n,), -3, dtype=torch.int16, device="cuda")
    d_y = torch.full((n,), -5, dtype=torch.int16, device="cuda")
    torch.cuda.synchronize()
    rc = lib.svt_sad_loop_kernel_b200_noalloc(None, n, d_blocks.data_ptr(), src_stride, src_step, d_ref.data_ptr() + ref_off0,
                                              raw * sub, ref_step, max(1, bh // sub), bw, raw, skip, sw, sh, d_best.data_ptr(),
                                              d_x.data_ptr(), d_y.data_ptr())
    assert rc == 0, lib.svtme_last_error()
    torch.cuda.synchronize()
    got = list(zip(d_best.cpu().numpy().astype(np.uint64).tolist(), d_x.cpu().numpy().tolist(), d_y.cpu().numpy().tolist()))
    for i in range(n):
        want = existing_entry_point(lib, blocks[i], src_stride, ref, ref_off0 + i * ref_step, raw * sub, max(1, bh // sub), bw, raw,
                                    skip, sw, sh)
        assert got[i] == want, ("search", i, (bw, bh, sw, sh, sub, skip), got[i], want)


def test_noalloc_sad_loop_equals_existing_entry_point_on_two_calls_in_a_row():
    import torch
    lib = U.product_lib()
    assert lib.svtme_noalloc_setup(torch.cuda.current_device()) == 0, lib.svtme_last_error()
    rng = np.random.default_rng(20261105)
    # first call: many searches (fills a long prefix of the static array); second call right after it: fewer searches,
    # other pixels and another geometry — stale keys of the first call must not leak into its outputs
    run_case(lib, torch, rng, 96, 16, 16, 24, 9, 2, 0)
    run_case(lib, torch, rng, 7, 32, 32, 8, 3, 2, 0)
    # HME-shaped blocks of every level, odd areas, the skip_search_line rule, full-height blocks
    for n, bw, bh, sw, sh, sub, skip in [(5, 16, 16, 16, 6, 2, 1), (3, 64, 64, 8, 3, 2, 0), (4, 24, 12, 7, 5, 1, 0), (2, 16, 8, 160, 3, 1, 1),
                                         (2, 8, 8, 1, 1, 1, 0)]:
        run_case(lib, torch, rng, n, bw, bh, sw, sh, sub, skip)


def test_noalloc_rejects_more_searches_than_the_static_array_holds_and_bad_arguments():
    import torch
    lib = U.product_lib()
    buf = torch.zeros(1 << 16, dtype=torch.uint8, device="cuda")
    out64 = torch.zeros(4, dtype=torch.int64, device="cuda")
    out16 = torch.zeros(8, dtype=torch.int16, device="cuda")
    args = (buf.data_ptr(), 64, 0, buf.data_ptr() + 4096, 256, 0, 8, 16, 256, 0, 8, 3, out64.data_ptr(), out16.data_ptr(),
            out16.data_ptr() + 8)
    assert lib.svt_sad_loop_kernel_b200_noalloc(None, MAX_SEARCHES + 1, *args) == -1  # SVTME_ERR_INVALID
    assert b"maximum 8192" in lib.svtme_last_error()
    assert lib.svt_sad_loop_kernel_b200_noalloc(None, 2, None, *args[1:]) == -1      # null input
    bad = list(args)
    bad[7] = 65                                                                       # block wider than 64
    assert lib.svt_sad_loop_kernel_b200_noalloc(None, 2, *bad) == -1
    assert lib.svt_sad_loop_kernel_b200_noalloc(None, 0, *args) == 0                  # nothing to do
    assert lib.svt_sad_loop_kernel_b200_noalloc(None, 2, *args) == 0                  # and the valid call still works
    torch.cuda.synchronize()
    assert out64[:2].cpu().tolist() == [0, 0]                                         # all-zero block vs all-zero window
