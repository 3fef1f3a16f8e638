"""CPU pins of the product-grid oracle (oracle/grid2d_oracle.cpp: atomic / diatomic DFTGrid::eval_Fxc restated).

The reference's own run-time identities for this path (/root/reference/src/atomic/main.cpp:412-440,810-812): the grid
reproduces the analytic overlap and kinetic matrices, i.e. Nel = tr(P S) and Ekin = tr(P T); plus the variational
identity H_ij = dExc/dP_ij that ties the restated Fock assembly (increment_lda / _gga / tau terms) to the restated
energy, and the spherical limit against the independent sadatom oracle."""
import numpy as np
import pytest

import oracle
from grid2d_cases import ATOMIC, atomic_oracle, atomic_matrices, atomic_density

LDA, PBE, R2SCAN = (1, 12), (101, 130), (497, 498)


@pytest.mark.parametrize("case", ["ne_l0", "ne_l2m2", "c_hip_l1m1"])
def test_grid_reproduces_overlap_and_kinetic(oracle_lib, case):
    o, lval, mval, rs = atomic_oracle(case)
    S, T, V = atomic_matrices(rs, lval, ATOMIC[case][0])
    P = atomic_density(case)
    H, exc, nel, ekin = o.eval_fxc(*R2SCAN, P)
    # the modified Gauss-Chebyshev radial rule is spectrally accurate, not exact, for these polynomials
    assert abs(nel - np.sum(P * S)) <= 2e-8 * abs(nel)
    assert abs(ekin - np.sum(P * T)) <= 2e-8 * abs(ekin)
    assert np.max(np.abs(H - H.T)) <= 1e-10 * np.max(np.abs(H))


@pytest.mark.parametrize("funcs", [LDA, PBE, R2SCAN])
def test_fock_matrix_is_energy_derivative(oracle_lib, funcs):
    o, lval, mval, rs = atomic_oracle("c_hip_l1m1")
    P = atomic_density("c_hip_l1m1")
    H, exc, nel, ekin = o.eval_fxc(*funcs, P)
    n = P.shape[0]
    nr = rs.nbf
    eps = 1e-5
    # a same-element diagonal pair, a same-element off-diagonal pair, a pair across angular blocks
    for (i, j) in [(2, 2), (1, 3), (2, nr + 3), (nr + 1, 2 * nr + 2)]:
        E = np.zeros_like(P)
        E[i, j] += 1.0
        E[j, i] += 1.0
        ep = o.eval_fxc(*funcs, P + eps * E)[1]
        em = o.eval_fxc(*funcs, P - eps * E)[1]
        fd = (ep - em) / (2 * eps)
        want = H[i, j] + H[j, i]
        assert abs(fd - want) <= 2e-6 * max(1.0, abs(want)), (funcs, i, j, fd, want)


def test_spherical_limit_matches_sadatom_oracle(oracle_lib):
    """A spherically averaged density (equal occupation of every m of a shell) in the atomic program gives the
    sadatom program's Exc / Nel and its radial Fock blocks (two independently restated loops)."""
    Z, primbas, nnodes, nelem, lmax, mmax, nocc, lang, mang = ATOMIC["ne_l2m2"]
    o, lval, mval, rs = atomic_oracle("ne_l2m2")
    so = oracle.sadatom_from_setup(rs, lmax, nL=1, exchange=False)
    nr = rs.nbf
    rng = np.random.RandomState(3)
    Pl = np.zeros((nr, nr, lmax + 1), order="F")
    for l in range(lmax + 1):
        c = rng.uniform(-1, 1, (nr, 2)) * np.exp(-0.4 * np.arange(nr))[:, None]
        Pl[:, :, l] = (2 * l + 1) * (c @ c.T)
    P = np.zeros((len(lval) * nr,) * 2, order="F")
    for a, l in enumerate(lval):
        P[a * nr:(a + 1) * nr, a * nr:(a + 1) * nr] = Pl[:, :, l] / (2 * l + 1)
    for funcs in (PBE, R2SCAN):
        H, exc, nel, ekin = o.eval_fxc(*funcs, P)
        Hs, excs, nels = so.eval_fxc(*funcs, Pl / (4 * np.pi))
        assert abs(nel - nels) <= 1e-10 * abs(nel)
        assert abs(exc - excs) <= 1e-10 * abs(exc)
        # sadatom's H is per l; the caller divides by 4 pi (sadatom/solver.cpp:806)
        for a, l in enumerate(lval):
            blk = H[a * nr:(a + 1) * nr, a * nr:(a + 1) * nr]
            assert np.max(np.abs(blk - Hs[:, :, l] / (4 * np.pi))) <= 1e-9 * max(1.0, np.max(np.abs(blk)))
