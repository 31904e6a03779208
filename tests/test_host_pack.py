"""Host side of the packed PCIe transfer (n2v2r_host_pack_nonzeros, host_pack.cpp): bit-exact round trip against a
NumPy expansion of the bitmap.  Pure host code: runs without a GPU."""
import ctypes as C

import numpy as np
import pytest

from node2vec2rank_b200 import _native as N


def _p(x):
    return C.c_void_p(x.ctypes.data)


def _pack(a, cols, capacity, threads):
    rows, ld = a.shape
    words = (cols + 31) // 32
    bitmap = np.zeros((rows, max(words, 1)), np.uint32)
    values = np.zeros(max(capacity, 1) + 8, np.float32)
    row_start = np.full(max(rows, 1), -7, np.int64)
    nnz = C.c_int64(-5)
    rc = N.load().n2v2r_host_pack_nonzeros(_p(a), rows, cols, ld, _p(bitmap), _p(values), capacity, _p(row_start), threads,
                                           C.byref(nnz))
    return rc, nnz.value, bitmap[:, :words], values, row_start


def _expand(bitmap, values, row_start, rows, cols):
    out = np.zeros((rows, cols), np.float32)
    if rows == 0 or cols == 0:
        return out
    bits = np.unpackbits(bitmap.view(np.uint8).reshape(rows, -1), axis=1, bitorder="little")[:, :cols].astype(bool)
    for r in range(rows):
        k = int(bits[r].sum())
        out[r, bits[r]] = values[row_start[r]:row_start[r] + k]
    return out


@pytest.mark.parametrize("rows,cols,pad,density,threads", [(37, 1000, 0, 0.1, 4), (64, 33, 3, 0.5, 3), (5, 31, 0, 1.0, 8),
                                                            (300, 257, 7, 0.02, 1), (9, 64, 0, 0.0, 2)])
def test_pack_round_trip_is_bit_exact(rows, cols, pad, density, threads):
    rng = np.random.default_rng(rows * 1000 + cols)
    a = np.zeros((rows, cols + pad), np.float32)
    body = rng.standard_normal((rows, cols)).astype(np.float32)
    body[rng.random((rows, cols)) >= density] = 0
    a[:, :cols] = body
    a[:, cols:] = 9.0                                   # padding beyond `cols` must be ignored
    if density > 0 and rows > 3 and cols > 40:
        a[1, 2] = -0.0                                  # all-bits test: -0.0 and NaN are values, not zeros
        a[2, 33] = np.nan
    want_nnz = int(np.count_nonzero(a[:, :cols].view(np.uint32)))
    rc, nnz, bitmap, values, row_start = _pack(a, cols, rows * cols, threads)
    assert rc == 0 and nnz == want_nnz
    got = _expand(bitmap, values, row_start, rows, cols)
    assert np.array_equal(got.view(np.uint32), np.ascontiguousarray(a[:, :cols]).view(np.uint32))


def test_pack_overflow_and_arguments():
    a = np.ones((40, 40), np.float32)
    rc, nnz, *_ = _pack(a, 40, 100, 4)
    assert rc == 0 and nnz == -1                         # does not fit: the caller copies the layer dense
    rc, nnz, *_ = _pack(a, 40, 1600, 4)
    assert rc == 0 and nnz == 1600
    rc, nnz, *_ = _pack(np.zeros((0, 8), np.float32), 8, 10, 2)
    assert rc == 0 and nnz == 0
    lib = N.load()
    n = C.c_int64(0)
    assert lib.n2v2r_host_pack_nonzeros(None, 4, 4, 4, None, None, 4, None, 1, C.byref(n)) == 1      # N2V2R_EINVAL
    assert b"bad argument" in lib.n2v2r_last_error()


def test_host_threads_are_shared_between_local_ranks(monkeypatch):
    from node2vec2rank_b200 import device as D
    monkeypatch.delenv("LOCAL_WORLD_SIZE", raising=False)
    cores = D.host_threads()
    assert cores >= 1
    monkeypatch.setenv("LOCAL_WORLD_SIZE", "4")
    assert D.host_threads() == max(1, cores // 4)
    monkeypatch.setenv("LOCAL_WORLD_SIZE", "not-a-number")
    assert D.host_threads() == cores
