"""CPU tests: the oracle's setup restatement and the product's host setup library against the
known-answer vectors the reference's own tests hold, and against each other."""
import ctypes
import json
import math
import os

import numpy as np
import pytest

import oracle
from oracle import fem_oracle as fo

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
c_double_p = ctypes.POINTER(ctypes.c_double)


def _p(a):
    return a.ctypes.data_as(c_double_p)


# Maple table of /root/reference/src/atomic/inttest.cpp:58-98 (one-sided integrals in units of R; the
# test's 4*pi is the angular factor that coulomb() applies later as Lfac, basis.cpp:307)
MAPLE = np.array([[47 / 180, 11 / 360, 11 / 360, 1 / 90], [1 / 10, 1 / 40, 1 / 40, 1 / 60], [1 / 10, 1 / 40, 1 / 40, 1 / 60],
                  [3 / 20, 7 / 120, 7 / 120, 1 / 15]])


@pytest.mark.parametrize("R", [1.0, 2.5])
def test_maple_table_oracle(R):
    xq, wq = fo.chebyshev(200)
    t = fo.twoe_integral(0.0, R, xq, wq, fo.PolyBasis(4, 2), 0)
    assert np.max(np.abs(t - (MAPLE + MAPLE.T) * R)) < 1e-14 * R


@pytest.mark.parametrize("R", [1.0, 2.5])
def test_maple_table_product_host(native, R):
    out = np.zeros((4, 4), order="F")
    assert native.hfh_twoe_integral(4, 2, 200, ctypes.c_double(0.0), ctypes.c_double(R), 0, _p(out)) == 0
    assert np.max(np.abs(out - (MAPLE + MAPLE.T) * R)) < 1e-14 * R


def test_inner_integral_closed_forms():
    # inttest.cpp:43-47: inner integrals of the two linear functions on [0,R]
    R, n = 2.0, 60
    xq, wq = fo.chebyshev(n)
    poly = fo.PolyBasis(4, 2)
    r = 0.5 * R + 0.5 * R * xq
    # reuse the oracle's twoe_integral internals: outer integral with delta weights is not exposed, so check
    # the full integral against the closed form of sum_p w_p B_i B_j(r_p) phi_kl(r_p) instead
    phi = np.stack([1 - r / R + r**2 / (3 * R * R), r * (3 * R - 2 * r) / (6 * R * R), r * (3 * R - 2 * r) / (6 * R * R),
                    r**2 / (3 * R * R)], axis=1)
    bf = poly.eval_dnf(xq, 0, 0.5 * R)
    bfprod = (bf[:, :, None] * bf[:, None, :]).reshape(n, 4) * (wq * 0.5 * R)[:, None]
    ints = bfprod.T @ phi
    t = fo.twoe_integral(0.0, R, xq, wq, poly, 0)
    assert np.max(np.abs(t - (ints + ints.T))) < 1e-13


def test_gaunt_golden_table(native):
    g = json.load(open(os.path.join(GOLD, "gaunt_m0.json")))
    assert len(g) == 550
    eps = np.finfo(float).eps
    for e in g:
        a, ref = e["args"], e["ref"]
        if e["modified"]:
            assert abs(native.hfh_modified_gaunt(*a) - ref) < 4 * eps * (1 + abs(ref))
        else:
            assert abs(native.hfh_gaunt(*a) - ref) < 4 * eps * (1 + abs(ref))
            assert abs(oracle.gaunt(*a) - ref) < 4 * eps * (1 + abs(ref))


def test_gaunt_m_nonzero_against_sympy(native):
    from sympy.physics.wigner import gaunt as sgaunt
    for (L, M, l, m, lp, mp) in [(2, 1, 1, 0, 1, 1), (3, -2, 2, -1, 1, -1), (4, 2, 2, 1, 2, 1), (6, 0, 3, 2, 3, -2), (5, 3, 3, 1, 4, 2)]:
        # int conj(Y_LM) Y_lm Y_l'm' = (-1)^M int Y_{L,-M} Y_lm Y_l'm'
        ref = float(sgaunt(L, l, lp, -M, m, mp)) * (-1) ** M
        assert abs(native.hfh_gaunt(L, M, l, m, lp, mp) - ref) < 1e-14
        assert abs(oracle.gaunt(L, M, l, m, lp, mp) - ref) < 1e-14


def test_lobatto_nodes_match_reference_table(native):
    lob = json.load(open(os.path.join(GOLD, "lobatto_nodes.json")))
    for key, ref in lob.items():
        n = int(key)
        x = np.zeros(n)
        assert native.hfh_lobatto_nodes(n, _p(x)) == 0
        tol = 2e-14 if n == 12 else 0.0  # the reference's n=12 table deviates from the true roots in the 15th digit
        assert np.max(np.abs(x - np.array(ref))) <= tol
        assert np.max(np.abs(fo.lobatto_nodes(n) - np.array(ref))) <= tol


def test_chebyshev_rule_integrates_polynomials(native):
    for n in (10, 75):
        x, w = fo.chebyshev(n)
        xp, wp = np.zeros(n), np.zeros(n)
        assert native.hfh_chebyshev(n, _p(xp), _p(wp)) == 0
        assert np.array_equal(x, xp) and np.array_equal(w, wp)
        # mapped (non-Gaussian) rule: spectrally convergent, not exact for polynomials
        tol = 1e-6 if n == 10 else 1e-12
        assert abs(np.sum(w) - 2.0) < tol
        assert abs(np.sum(w * x**2) - 2.0 / 3.0) < tol


@pytest.mark.parametrize("primbas,nel,taylor", [(4, 5, -1), (5, 4, 20)])
def test_host_setup_matches_oracle_setup(native, primbas, nel, taylor):
    import helfem_b200 as hb
    b = hb.SadatomBasis(Z=10, nelem=nel, nnodes=15, primbas=primbas, taylor_order=taylor, lmax=2)
    rs = fo.RadialSetup(primbas, 15, fo.get_grid(40.0, nel), taylor_order=taylor)
    assert b.nbf == rs.nbf and list(b.first) == rs.first and list(b.last) == rs.last
    assert b.taylor_cutoff == rs.cutoff
    for e in range(nel):
        r, w, bf, df, lf = b.tables(e)
        for x, y in ((r, rs.r(e)), (w, rs.wrad(e)), (bf, rs.get_bf(e)), (df, rs.get_df(e)), (lf, rs.get_lf(e))):
            assert np.max(np.abs(x - y)) <= 1e-13 * np.max(np.abs(y))
    for A, B in ((b.overlap(), rs.overlap()), (b.kinetic(), rs.kinetic()), (b.kinetic_l(), rs.kinetic_l()), (b.nuclear(), rs.nuclear(10))):
        assert np.max(np.abs(A - B)) <= 1e-13 * np.max(np.abs(B))
    b.compute_tei(3, True)
    for L in (0, 2):
        for e in (0, 1, nel - 1):
            a, bb, t, k = b.tei(L, e)
            t2 = rs.twoe_integral(L, e)
            n = rs.basis(e).nbf
            assert np.max(np.abs(t - t2)) <= 1e-13 * np.max(np.abs(t2))
            assert np.max(np.abs(k - fo.exchange_tei(t2, n, n, n, n))) <= 1e-13 * np.max(np.abs(t2))
            assert np.max(np.abs(a - rs.radial_integral(L, e))) <= 1e-13 * np.max(np.abs(a))
            assert np.max(np.abs(bb - rs.radial_integral(-L - 1, e))) <= 1e-13 * np.max(np.abs(bb))


def test_tei_permutational_symmetry():
    # utils.cpp:96-128 (check_tei_symmetry): (ij|kl) = (ji|kl) = (ij|lk) = (kl|ij)
    rs = fo.RadialSetup(4, 6, fo.get_grid(10.0, 3))
    t = rs.twoe_integral(1, 1)
    n = rs.basis(1).nbf
    t4 = t.reshape((n, n, n, n), order="F")
    assert np.max(np.abs(t4 - t4.transpose(1, 0, 2, 3))) < 1e-15
    assert np.max(np.abs(t4 - t4.transpose(0, 1, 3, 2))) < 1e-15
    assert np.max(np.abs(t4 - t4.transpose(2, 3, 0, 1))) < 1e-14 * np.max(np.abs(t4))


def test_hydrogen_ground_state_from_setup_matrices():
    # end-to-end pin of S, T, Vnuc: the 1s level of hydrogen is -1/2 Eh
    import scipy.linalg
    rs = fo.RadialSetup(4, 15, fo.get_grid(40.0, 5))
    E = scipy.linalg.eigh(rs.kinetic() + rs.nuclear(1), rs.overlap(), eigvals_only=True)
    assert abs(E[0] + 0.5) < 1e-9
    assert abs(E[1] + 0.125) < 1e-8
