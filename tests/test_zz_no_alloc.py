"""ohcu_dsp_idct_noalloc — the idct[] entry with its device staging in a static array (no cudaMalloc): same
results as ohcu_dsp_idct, bit for bit (integer path, tolerance 0), on two calls in a row that share the array;
sizes beyond the array are refused.  Collected last on purpose: the file name sorts
after every other test file.  Runs alone in a fresh process as well as after the others."""
import ctypes as C

import numpy as np
import pytest

from hevc_b200 import abi

pytestmark = pytest.mark.gpu


def _blocks(rng, log2, nblocks, col_limit):
    """coefficients at the positions the reference's residual parser can fill for this col_limit"""
    n = 1 << log2
    c = np.zeros((nblocks, n, n), np.int16)
    lim = min(col_limit, n)
    c[:, :lim, :lim] = rng.integers(-2000, 2001, size=(nblocks, lim, lim))
    c[:, 0, 0] = rng.integers(-32768, 32768, size=nblocks)
    return c


def _reference(lib, bd, blocks, log2, col_limit):
    out = blocks.copy()
    for b in out:
        assert lib.ohcu_dsp_idct(bd, b.ctypes.data, log2, col_limit) == 0, lib.ohcu_last_error()
    return out


def test_idct_noalloc_matches_idct_on_two_calls_in_a_row():
    lib = abi.load()
    rng = np.random.default_rng(2026)
    checked = 0
    for bd in (8, 10, 12):
        for log2, col_limit in ((2, 4), (3, 8), (3, 4), (4, 16), (4, 8), (5, 32), (5, 12), (5, 4)):
            n = 1 << log2
            cap = abi.NOALLOC_MAX_BYTES // (2 * n * n)
            # first call fills the whole array, the second uses a part of it over the first call's leftovers
            for nblocks in (cap, max(1, cap // 3)):
                blocks = _blocks(rng, log2, nblocks, col_limit)
                want = _reference(lib, bd, blocks, log2, col_limit)
                got = blocks.copy()
                rc = lib.ohcu_dsp_idct_noalloc(bd, got.ctypes.data, log2, col_limit, nblocks)
                assert rc == 0, lib.ohcu_last_error()
                assert np.array_equal(got, want), f"bd {bd} {n}x{n} col_limit {col_limit} x{nblocks}: " \
                                                  f"{np.argwhere(got != want)[:4].tolist()}"
                assert not np.array_equal(got, blocks)
                checked += nblocks
    assert checked > 1000


def test_idct_noalloc_rejects_sizes_beyond_the_static_array():
    lib = abi.load()
    buf = np.zeros(abi.NOALLOC_MAX_BYTES // 2 + 32 * 32, np.int16)
    assert lib.ohcu_dsp_idct_noalloc(10, buf.ctypes.data, 5, 32, abi.NOALLOC_MAX_BYTES // 2048 + 1) == -1
    msg = lib.ohcu_last_error()
    assert b"OHCU_NOALLOC_MAX_BYTES" in msg and b"65536" in msg, msg
    assert lib.ohcu_dsp_idct_noalloc(10, buf.ctypes.data, 6, 32, 1) == -1
    assert lib.ohcu_dsp_idct_noalloc(11, buf.ctypes.data, 5, 32, 1) == -1
    assert lib.ohcu_dsp_idct_noalloc(10, None, 5, 32, 1) == -1
    # the limit itself is accepted
    assert lib.ohcu_dsp_idct_noalloc(10, buf.ctypes.data, 5, 32, abi.NOALLOC_MAX_BYTES // 2048) == 0, lib.ohcu_last_error()
