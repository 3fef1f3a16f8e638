"""GPU tests of the family-sharded atomic / diatomic operators (run with -m gpu): on ONE device, build
nranks contexts with hfb200_set_shard(rank, nranks); every partial must be non-trivial and the partials
must add up to the oracle's full J / K (<= 1e-9).  The real NCCL all-reduce of the same partials is
exercised by tools/shard_nccl_check.py under torchrun (gpurun --gpus 2)."""
import numpy as np
import pytest

import oracle
from oracle import fem_oracle as fo
from util import splitmix64

pytestmark = pytest.mark.gpu


def tol(ref):
    return max(1e-9, 5e-14 * float(np.max(np.abs(ref))))


@pytest.mark.parametrize("nranks", [2, 3])
def test_atomic_family_shards_sum_to_full(native, nranks):
    import helfem_b200 as hb
    Z, primbas, nnodes, nelem, lmax, mmax = 10, 4, 8, 3, 2, 2
    rs = fo.RadialSetup(primbas, nnodes, fo.get_grid(40.0, nelem), taylor_order=-1)
    nL = 2 * lmax + 1
    dL, dm1L, tei, ktei = [], [], [], []
    for L in range(nL):
        for e in range(rs.nel):
            dL.append(rs.radial_integral(L, e))
            dm1L.append(rs.radial_integral(-L - 1, e))
            t = rs.twoe_integral(L, e)
            tei.append(t)
            n = rs.basis(e).nbf
            ktei.append(fo.exchange_tei(t, n, n, n, n))
    Jsum = Ksum = None
    for rank in range(nranks):
        b = hb.AtomicBasis(Z=Z, nelem=nelem, nnodes=nnodes, primbas=primbas, lmax=lmax, mmax=mmax)
        b.set_shard(rank, nranks)
        b.compute_tei(True)
        if rank == 0:
            o = oracle.AtomicOracle(b.lval, b.mval, rs.first, rs.last, dL, dm1L, tei, ktei)
            n = b.nbf
            R = splitmix64(n * n).reshape(n, n)
            P = np.asfortranarray(0.5 * (R + R.T))
            Jo, Ko = o.coulomb(P), o.exchange(P)
        J, K = b.coulomb(P), b.exchange(P)
        assert np.max(np.abs(J)) > 0 and np.max(np.abs(K)) > 0
        assert np.max(np.abs(J - Jo)) > 1e-6 and np.max(np.abs(K - Ko)) > 1e-6  # a partial, not the whole
        Jsum = J if Jsum is None else Jsum + J
        Ksum = K if Ksum is None else Ksum + K
    assert np.max(np.abs(Jsum - Jo)) <= tol(Jo)
    assert np.max(np.abs(Ksum - Ko)) <= tol(Ko)


@pytest.mark.parametrize("nranks", [2, 4])
def test_diatomic_family_shards_sum_to_full(native, nranks):
    import helfem_b200 as hb
    args = dict(nelem=2, nnodes=6, primbas=4, lmax=2, mmax=1)
    Jsum = Ksum = None
    for rank in range(nranks):
        b = hb.DiatomicBasis(7, 7, 2.068, **args)
        b.set_shard(rank, nranks)
        b.compute_tei(True)
        if rank == 0:
            dis = [[b.tei(w, r, e) for r in range(b.nlm) for e in range(b.nel)] for w in range(4)]
            tei = [b.tei(4 + ab, r, e) for r in range(b.nlm) for e in range(b.nel) for ab in range(4)]
            ktei = [b.tei(8 + ab, r, e) for r in range(b.nlm) for e in range(b.nel) for ab in range(4)]
            o = oracle.DiatomicOracle(b.lval, b.mval, b.first, b.last, b.Lmap, b.Mmap, b.Rhalf, dis, tei, ktei,
                                      b.pure_indices())
            n = b.nbf
            R = splitmix64(n * n).reshape(n, n)
            C = splitmix64(n * 4, seed=7).reshape(n, 4) * np.exp(-0.05 * np.arange(n))[:, None]
            P = np.asfortranarray(C @ C.T + 1e-2 * 0.5 * (R + R.T))
            Jo, Ko = o.coulomb(P), o.exchange(P)
        J, K = b.coulomb(P), b.exchange(P)
        assert np.max(np.abs(J)) > 0 and np.max(np.abs(K)) > 0
        Jsum = J if Jsum is None else Jsum + J
        Ksum = K if Ksum is None else Ksum + K
    assert np.max(np.abs(Jsum - Jo)) <= tol(Jo)
    assert np.max(np.abs(Ksum - Ko)) <= tol(Ko)


def test_set_shard_after_tei_is_refused(native):
    import helfem_b200 as hb
    b = hb.AtomicBasis(Z=2, nelem=2, nnodes=6, lmax=0, mmax=0)
    b.compute_tei(False)
    n = b.nbf
    b.coulomb(np.asfortranarray(np.eye(n)))  # creates the context and uploads the integrals
    import ctypes
    native.hfb200_set_shard.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
    assert native.hfb200_set_shard(None, 0, 1) == hb.EINVAL           # null context
    assert native.hfb200_set_shard(b.context(), 0, 2) == hb.EINVAL    # the context already holds the full set of integrals
    assert b"precede" in hb.lib().hfb200_last_error()
    with pytest.raises(Exception):                                    # the host class refuses as well (std::logic_error)
        b.set_shard(0, 2)
    own = np.zeros(1, dtype=np.int32)
    native.hfb200_shard_owner.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    assert native.hfb200_shard_owner(b.context(), own.ctypes.data_as(ctypes.c_void_p), 1) == 0 and own[0] == 0
