"""CPU tests of the host side of the dense-matrix operator boundary (helfem_b200/host/blockscan.cpp): the scan of P for its
non-zero angular blocks and the non-temporal zero fill of the result are product code that runs on the host inside every
coulomb / exchange / eval_Fxc call (reference boundary: dense arma::mat in and out, diatomic/basis.cpp:1359,1532)."""
import ctypes

import numpy as np
import pytest

from util import splitmix64


def _lib():
    import helfem_b200
    L = helfem_b200.lib()
    L.hfh_scan_nonzero_blocks.restype = None
    L.hfh_zero_fill.restype = None
    return L


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


@pytest.mark.parametrize("nang,nrad,ragged", [(1, 1, False), (5, 7, False), (13, 10, True), (40, 3, True)])
def test_scan_matches_numpy(nang, nrad, ragged):
    # ragged: the first radial function of some angular functions is absent (the diatomic pure basis)
    skip = (np.arange(nang) % 3 == 1).astype(np.int32) if ragged else np.zeros(nang, dtype=np.int32)
    cnt = (nrad - skip).astype(np.int32)
    r0 = np.concatenate([[0], np.cumsum(cnt)[:-1]]).astype(np.int32)
    n = int(cnt.sum())
    ld = n + 3
    M = np.zeros((ld, n), order="F")
    want = np.zeros((nang, nang), dtype=np.uint8)
    rnd = splitmix64(4 * nang * nang, seed=5)
    specials = [1.0, -1e-300, 5e-324, np.nan, np.inf]
    for i in range(nang):
        for l in range(nang):
            if cnt[i] == 0 or cnt[l] == 0:
                continue
            u = rnd[4 * (i * nang + l)]
            if u < 0.35:   # one entry somewhere in the block, including its last row / column
                a = int((rnd[4 * (i * nang + l) + 1] + 1) * 0.5 * cnt[i]) % cnt[i]
                b = int((rnd[4 * (i * nang + l) + 2] + 1) * 0.5 * cnt[l]) % cnt[l]
                M[r0[i] + a, r0[l] + b] = specials[(i + 2 * l) % len(specials)]
                want[i, l] = 1
            elif u < 0.45:  # signed zeros only: still a zero block
                M[r0[i]:r0[i] + cnt[i], r0[l]:r0[l] + cnt[l]] = -0.0
    M[n:, :] = 7.0  # the padding rows below the matrix must not be looked at
    nz = np.full(nang * nang, 9, dtype=np.uint8)
    _lib().hfh_scan_nonzero_blocks(_p(M), ld, nang, _p(r0), _p(cnt), _p(nz))
    assert np.array_equal(nz.reshape(nang, nang), want)


@pytest.mark.parametrize("n,ld", [(1, 1), (5, 8), (131, 140), (1000, 1003)])
def test_zero_fill_touches_only_the_matrix(n, ld):
    M = np.full((ld, n + 2), 3.0, order="F")
    _lib().hfh_zero_fill(_p(M), ld, n)
    assert not M[:n, :n].any()
    assert np.all(M[n:, :] == 3.0) and np.all(M[:, n:] == 3.0)
    off = np.full(ld * n + 5, 2.0)   # an unaligned start address
    v = off[1:]
    _lib().hfh_zero_fill(_p(v), ld, n)
    A = v[:ld * n].reshape(n, ld).T
    assert not A[:n, :n].any() and off[0] == 2.0 and np.all(A[n:, :] == 2.0)
