"""b200cs_host_copy (csrc/host_stage.cu): the all-core copy of a caller's pageable array into a pinned
staging slab.  Host code only, so it runs without a GPU: sizes around the part and vector boundaries,
unaligned sources and destinations, repeated calls on the persistent pool, a forked child."""
import ctypes
import os

import numpy as np
import pytest

from fv3net_b200 import _lib


@pytest.fixture(scope="module")
def lib():
    return _lib.load()


def _copy(lib, dst, src, nthreads):
    assert lib.b200cs_host_copy(dst.ctypes.data, src.ctypes.data, src.nbytes, nthreads) == 0


@pytest.mark.parametrize("nbytes", [0, 1, 63, 255, 256, 4097, (1 << 20) - 1, (1 << 20) + 1, 5 * (1 << 20) + 77])
@pytest.mark.parametrize("nthreads", [1, 3, 8])
def test_host_copy_sizes_and_alignment(lib, nbytes, nthreads):
    rng = np.random.default_rng(nbytes + nthreads)
    for so, do in [(0, 0), (1, 0), (0, 3), (5, 9)]:
        src_buf = rng.integers(0, 256, size=nbytes + 64, dtype=np.uint8)
        dst_buf = np.full(nbytes + 128, 0xAB, dtype=np.uint8)
        src, dst = src_buf[so:so + nbytes], dst_buf[32 + do:32 + do + nbytes]
        _copy(lib, dst, src, nthreads)
        assert np.array_equal(dst, src)
        # nothing outside the destination was touched
        assert (dst_buf[:32 + do] == 0xAB).all() and (dst_buf[32 + do + nbytes:] == 0xAB).all()


def test_host_copy_pool_is_reusable_and_survives_fork(lib):
    rng = np.random.default_rng(3)
    src = rng.random(3 << 18)  # 6 MB
    dst = np.empty_like(src)
    for nthreads in (8, 2, 16, 8):  # the pool grows on demand and fewer workers than it holds may be asked for
        dst[:] = 0
        _copy(lib, dst, src, nthreads)
        assert np.array_equal(dst, src)
    pid = os.fork()
    if pid == 0:  # the child has none of the parent's workers
        ok = 1
        try:
            d2 = np.zeros_like(src)
            _copy(lib, d2, src, 8)
            ok = 0 if np.array_equal(d2, src) else 1
        finally:
            os._exit(ok)
    _, status = os.waitpid(pid, 0)
    assert os.WIFEXITED(status) and os.WEXITSTATUS(status) == 0
    dst[:] = 0
    _copy(lib, dst, src, 8)
    assert np.array_equal(dst, src)
