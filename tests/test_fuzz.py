"""Differential tests on random stage configurations (tests/_fuzz.py). CPU: the oracle restatement against the compiled
reference (needs oracle/_ref). GPU (-m gpu): the CUDA path against the CPU checker. tools/fuzz_oracle_vs_ref.py and
tools/fuzz_gpu.py run the same generator for as many configurations as wanted (round 2: 220 on the CPU and, after the fix, 704 on the GPU without a mismatch; on
the GPU the first 36 found that an incomplete b64 at the right edge — 1/16 block narrower than 16 — must not skip the even
pre-HME search rows under prehme_skip_search_line, compute_sad_c.c:74-79)."""
import pytest

import _fuzz as F
import _util as U


@pytest.mark.skipif(U.ref_lib() is None, reason="oracle/_ref not built (no /root/reference here)")
def test_oracle_equals_compiled_reference_on_random_configurations():
    assert F.run_cpu(11, 6) == []


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [5, 6])
def test_gpu_equals_cpu_checker_on_random_configurations(seed):
    assert F.run_gpu(seed, 36) == []
