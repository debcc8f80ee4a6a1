"""Host code of the host-streamed route (no GPU): `scvi_host_gather_rows` (persistent host thread pool) against numpy
fancy indexing -- the reference's `data[indexes]` of src/scvi/data/_anntorchdataset.py:143-152."""
import ctypes as C

import numpy as np
import pytest

from scvi_tools_b200._lib import load


@pytest.mark.parametrize("dtype,G,pad", [(np.uint8, 2000, 0), (np.uint16, 333, 5), (np.float32, 100, 0), (np.uint8, 1, 0)])
@pytest.mark.parametrize("threads", [1, 4, 0])
def test_host_gather_rows_matches_numpy(dtype, G, pad, threads):
    lib = load()
    rng = np.random.default_rng(0)
    n = 5000
    Xs = rng.integers(0, 200, size=(n, G + pad)).astype(dtype)
    X = Xs[:, :G]  # strided rows when pad > 0
    for B in (0, 1, 37, 4096):
        idx = rng.integers(0, n, size=B).astype(np.int64)
        out = np.full((max(B, 1), G + 3), 7, dtype=dtype)
        rc = lib.scvi_host_gather_rows(C.c_void_p(X.ctypes.data), n, G * X.itemsize, X.strides[0], C.c_void_p(idx.ctypes.data),
                                       B, C.c_void_p(out.ctypes.data), out.strides[0], threads)
        assert rc == 0, lib.scvi_last_error_string().decode()
        assert np.array_equal(out[:B, :G], X[idx])
        assert (out[:, G:] == 7).all()  # nothing written beyond the row


def test_host_gather_rows_rejects_bad_index():
    lib = load()
    X = np.zeros((10, 16), dtype=np.uint8)
    idx = np.array([0, 10], dtype=np.int64)
    out = np.zeros((2, 16), dtype=np.uint8)
    rc = lib.scvi_host_gather_rows(C.c_void_p(X.ctypes.data), 10, 16, 16, C.c_void_p(idx.ctypes.data), 2,
                                   C.c_void_p(out.ctypes.data), 16, 2)
    assert rc != 0 and "out of range" in lib.scvi_last_error_string().decode()
