"""CPU pins of the diatomic setup chain (the tables the GPU path consumes):
  * Legendre P_L^M, Q_L^M for xi > 1 against mpmath (40 digits), definition by derivatives
  * the reference's own identity test, /root/reference/src/legendre/legendre_test.cpp:52-96: the Neumann
    expansion  4 pi/Rh sum_LM (-1)^M (L-|M|)!/(L+|M|)! P_L^|M|(xi<) Q_L^|M|(xi>) Y_LM(1) Y*_LM(2) = 1/r12,
    which pins the P/Q/Ylm/LMfac conventions of diatomic coulomb/exchange (basis.cpp:1421)
  * in-element two-electron blocks and P/Q disjoint integrals of the product host library against an
    independent Gauss-Legendre double quadrature over the two triangles mu2 < mu1, mu2 > mu1 with the
    oracle's own numpy shape functions."""
import ctypes

import numpy as np
import pytest

from oracle import fem_oracle as fo

dp = ctypes.POINTER(ctypes.c_double)


def plq(native, Lmax, Mmax, mu):
    native.hfh_plm_array.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_double, dp]
    native.hfh_qlm_array.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_double, dp]
    native.hfh_plm_array.restype = None
    native.hfh_qlm_array.restype = None
    P = np.zeros((Mmax + 1, Lmax + 1))
    Q = np.zeros((Mmax + 1, Lmax + 1))
    native.hfh_plm_array(Lmax, Mmax, float(mu), P.ctypes.data_as(dp))
    native.hfh_qlm_array(Lmax, Mmax, float(mu), Q.ctypes.data_as(dp))
    return P, Q


def test_legendre_against_mpmath(native):
    mpmath = pytest.importorskip("mpmath")
    mpmath.mp.dps = 40
    for mu in (1e-3, 0.05, 0.7, 2.5, 4.4):
        P, Q = plq(native, 16, 5, mu)
        x = mpmath.cosh(mu)
        for l, m in ((0, 0), (1, 0), (1, 1), (4, 2), (9, 5), (16, 0), (16, 5), (5, 5)):
            pr = mpmath.legenp(l, m, x, type=3)
            qr = mpmath.re((x * x - 1) ** (mpmath.mpf(m) / 2) * mpmath.diff(lambda z: mpmath.legenq(l, 0, z, type=3), x, m))
            assert abs((P[m, l] - pr) / pr) < 1e-13
            assert abs((Q[m, l] - qr) / qr) < 1e-12


def test_neumann_expansion_is_coulomb_kernel(native):
    from scipy.special import sph_harm_y
    Rh, Lmax = 0.2, 50
    # first pair = the reference's own test point (legendre_test.cpp:43-55); the series converges like
    # exp(-(mu> - mu<) L), so the other pairs keep a gap >= 1 in mu
    pts = [((0.1, 0.3 * np.pi, 0.125 * np.pi), (2.4, 0.7 * np.pi, 1.875 * np.pi)), ((2.0, 1.0, 0.0), (0.6, 2.5, 4.0)),
           ((0.3, 0.1, 1.0), (1.5, 3.0, 5.5))]

    def cart(mu, nu, phi):
        return np.array([Rh * np.sinh(mu) * np.sin(nu) * np.cos(phi), Rh * np.sinh(mu) * np.sin(nu) * np.sin(phi),
                         Rh * np.cosh(mu) * np.cos(nu)])

    for (a, b) in pts:
        lo, hi = (a, b) if a[0] < b[0] else (b, a)
        P, _ = plq(native, Lmax, Lmax, lo[0])
        _, Q = plq(native, Lmax, Lmax, hi[0])
        s = 0.0
        for L in range(Lmax + 1):
            for M in range(-L, L + 1):
                am = abs(M)
                ratio = 1.0
                for p in range(L - am + 1, L + am + 1):
                    ratio *= p
                y1 = sph_harm_y(L, M, a[1], a[2])
                y2 = sph_harm_y(L, M, b[1], b[2])
                s += (-1) ** M / ratio * P[am, L] * Q[am, L] * (y1 * np.conj(y2)).real
        s *= 4 * np.pi / Rh
        exact = 1.0 / np.linalg.norm(cart(*a) - cart(*b))
        assert abs(s - exact) < 1e-10 * exact, (s, exact)


def _gl(n, a, b):
    x, w = np.polynomial.legendre.leggauss(n)
    return 0.5 * (b - a) * x + 0.5 * (a + b), 0.5 * (b - a) * w


@pytest.mark.parametrize("L,M", [(0, 0), (2, 1), (3, 0)])
def test_inelement_tei_against_independent_quadrature(native, L, M):
    import helfem_b200 as hb
    nnodes = 4
    # a fine rule so that the reference's (non-Gaussian) Chebyshev scheme is converged well below the gate
    b = hb.DiatomicBasis(1, 1, 1.4, nelem=2, nnodes=nnodes, primbas=4, lmax=1, mmax=1, rmax=10.0, nquad=100)
    b.compute_tei(False)
    ilm = [i for i in range(b.nlm) if (b.Lmap[i], b.Mmap[i]) == (L, M)][0]
    import scipy.special as sp  # noqa: F401
    mumax = np.arccosh(10.0 / 0.7)
    bval = fo.get_grid(mumax, 2)
    iel = 1  # last element: LIP functions 0..n-2 (last one dropped by the boundary condition)
    lo, hi = bval[iel], bval[iel + 1]
    poly = fo.PolyBasis(4, nnodes)
    poly.drop_last(True, True)
    mid, half = 0.5 * (lo + hi), 0.5 * (hi - lo)

    def bf(mu):
        return poly.eval_dnf((np.atleast_1d(mu) - mid) / half, 0, half)

    def PQ(mu):
        P, Q = plq(native, L + 1, max(M, 1), float(mu))
        return P[M, L], Q[M, L]

    npb = bf(np.array([mid])).shape[1]
    ng = 30
    xo, wo = _gl(ng, lo, hi)
    ref = {ab: np.zeros((npb * npb, npb * npb)) for ab in range(4)}
    refP = {k: np.zeros((npb, npb)) for k in (0, 2)}
    refQ = {k: np.zeros((npb, npb)) for k in (0, 2)}
    for p, (m1, w1) in enumerate(zip(xo, wo)):
        B1 = bf(m1)[0]
        P1, Q1 = PQ(m1)
        pr1 = np.outer(B1, B1).ravel(order="F")
        for k in (0, 2):
            refP[k] += w1 * np.sinh(m1) * np.cosh(m1) ** k * P1 * np.outer(B1, B1)
            refQ[k] += w1 * np.sinh(m1) * np.cosh(m1) ** k * Q1 * np.outer(B1, B1)
        for (a2, b2, big) in ((lo, m1, True), (m1, hi, False)):   # mu2 < mu1: Q(mu1) P(mu2); mu2 > mu1: P(mu1) Q(mu2)
            xi, wi = _gl(ng, a2, b2)
            for m2, w2 in zip(xi, wi):
                B2 = bf(m2)[0]
                P2, Q2 = PQ(m2)
                pr2 = np.outer(B2, B2).ravel(order="F")
                rad = (Q1 * P2) if big else (P1 * Q2)
                base = w1 * w2 * np.sinh(m1) * np.sinh(m2) * rad
                for a in range(2):
                    for bb in range(2):
                        ref[2 * a + bb] += base * np.cosh(m1) ** (2 * a) * np.cosh(m2) ** (2 * bb) * np.outer(pr1, pr2)
    for ab in range(4):
        got = b.tei(4 + ab, ilm, iel)
        assert np.max(np.abs(got - ref[ab])) < 2e-7 * np.max(np.abs(ref[ab])), (ab, np.max(np.abs(got - ref[ab])), np.max(np.abs(ref[ab])))
    for w, (tab, k) in enumerate(((refP, 0), (refP, 2), (refQ, 0), (refQ, 2))):
        got = b.tei(w, ilm, iel)
        assert np.max(np.abs(got - tab[k])) < 1e-9 * np.max(np.abs(tab[k]))
