"""GPU parity tests of the `atomic` program's Gaunt-coupled coulomb / exchange (run with -m gpu):
CUDA path through the C ABI / host mirror class vs the CPU oracle (oracle/chan_oracle.cpp, the reference's
loop structure) on the same seeded densities.  The oracle's stored integrals come from the numpy setup
(oracle/fem_oracle.py), the product's from its own C++ host library.  Gate: max |dF| <= 1e-9."""
import numpy as np
import pytest
import scipy.linalg

import oracle
from oracle import fem_oracle as fo
from util import splitmix64

pytestmark = pytest.mark.gpu

CASES = {
    # name: (Z, primbas, nnodes, nelem, lmax, mmax, nocc)
    "ne_cfg1": (10, 4, 15, 10, 0, 0, 5),   # BASELINE configs[0]: Ne RHF, LIP-15, 10 elements, lmax=0
    "ne_l2m2": (10, 4, 10, 4, 2, 2, 5),
    "c_hip_l1m1": (6, 5, 6, 3, 1, 1, 3),
    "ar_l3m1": (18, 4, 8, 3, 3, 1, 9),
}
_cache = {}


def setup_case(name):
    if name not in _cache:
        import helfem_b200 as hb
        Z, primbas, nnodes, nelem, lmax, mmax, nocc = CASES[name]
        b = hb.AtomicBasis(Z=Z, nelem=nelem, nnodes=nnodes, primbas=primbas, lmax=lmax, mmax=mmax)
        b.compute_tei(True)
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
        o = oracle.AtomicOracle(b.lval, b.mval, rs.first, rs.last, dL, dm1L, tei, ktei)
        # density: lowest core-Hamiltonian orbitals of the product basis + a seeded dense symmetric perturbation
        # that populates every angular block (all M channels couple)
        S, H = b.overlap(), b.kinetic() + b.nuclear()
        E, C = scipy.linalg.eigh(H, S)
        P = 2.0 * C[:, :nocc] @ C[:, :nocc].T
        n = b.nbf
        R = splitmix64(n * n).reshape(n, n)
        P = np.asfortranarray(P + 1e-3 * np.max(np.abs(P)) * 0.5 * (R + R.T))
        _cache[name] = (b, o, P)
    return _cache[name]


def tol(ref):
    return max(1e-9, 5e-14 * float(np.max(np.abs(ref))))


@pytest.mark.parametrize("case", list(CASES))
def test_atomic_coulomb_parity(native, case):
    b, o, P = setup_case(case)
    J, Jo = b.coulomb(P), o.coulomb(P)
    d = np.max(np.abs(J - Jo))
    assert d <= tol(Jo), (d, np.max(np.abs(Jo)))
    assert np.max(np.abs(J - J.T)) <= 1e-9          # symmetric P -> symmetric J
    assert np.max(np.abs(b.coulomb(2.5 * P) - 2.5 * J)) <= 1e-8 * max(1.0, np.max(np.abs(J)))  # linearity


@pytest.mark.parametrize("case", list(CASES))
def test_atomic_exchange_parity(native, case):
    b, o, P = setup_case(case)
    K, Ko = b.exchange(P), o.exchange(P)
    d = np.max(np.abs(K - Ko))
    assert d <= tol(Ko), (d, np.max(np.abs(Ko)))
    assert np.max(np.abs(K - K.T)) <= 1e-9
    assert np.sum(K * P) < 0  # exchange() returns -K


def test_atomic_hartree_energy_of_closed_shell_is_positive_and_exchange_cancels_self_interaction(native):
    """One doubly occupied orbital: E_J = 1/2 tr(PJ) and E_K = 1/4 tr(P Ka-like) -> J - K/2 self-interaction
    identity for a two-electron closed shell: tr(P K)/2 = -tr(P J)/2 with K = exchange(P/2)*2."""
    b, o, _ = setup_case("ne_cfg1")
    S, H = b.overlap(), b.kinetic() + b.nuclear()
    E, C = scipy.linalg.eigh(H, S)
    Pa = np.asfortranarray(np.outer(C[:, 0], C[:, 0]))
    J = b.coulomb(Pa)
    Ka = b.exchange(Pa)
    # for a single spin-orbital the exchange energy cancels the Hartree self-repulsion exactly
    assert abs(np.sum(Pa * J) + np.sum(Pa * Ka)) <= 1e-9 * abs(np.sum(Pa * J))
    # hydrogen-like 1s with Z=10: J_1s1s = 5 Z / 8
    assert abs(np.sum(Pa * J) - 5 * 10 / 8) < 1e-6


def test_atomic_missing_tei_is_logic_error(native):
    import helfem_b200 as hb
    b = hb.AtomicBasis(Z=2, nelem=2, nnodes=6, lmax=0, mmax=0)
    with pytest.raises(ValueError):
        b.coulomb(np.zeros((b.nbf, b.nbf)))


def test_atomic_nonsymmetric_density(native):
    """coulomb / exchange are linear in an arbitrary (not necessarily symmetric) matrix, as in the reference."""
    b, o, P = setup_case("ne_l2m2")
    n = b.nbf
    A = np.asfortranarray(P + 0.2 * np.max(np.abs(P)) * np.triu(splitmix64(n * n, seed=78).reshape(n, n), 1))
    J, Jo = b.coulomb(A), o.coulomb(A)
    K, Ko = b.exchange(A), o.exchange(A)
    assert np.max(np.abs(J - Jo)) <= tol(Jo), np.max(np.abs(J - Jo))
    assert np.max(np.abs(K - Ko)) <= tol(Ko), np.max(np.abs(K - Ko))
