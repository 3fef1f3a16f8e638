"""Shared set-up of the product-grid (atomic / diatomic DFTGrid) parity cases: oracle objects and seeded densities."""
import numpy as np
import scipy.linalg

import oracle
from oracle import fem_oracle as fo
from util import splitmix64

ATOMIC = {
    # name: (Z, primbas, nnodes, nelem, lmax, mmax, nocc, lang, mang)    lang/mang default: 4 lmax + 10, 4 mmax + 5
    "ne_l0": (10, 4, 8, 4, 0, 0, 5, 10, 5),               # (atomic/main.cpp:393,401)
    "ne_l2m2": (10, 4, 6, 3, 2, 2, 5, 18, 13),
    "c_hip_l1m1": (6, 5, 4, 3, 1, 1, 3, 14, 9),
    "ar_l3m1": (18, 4, 6, 2, 3, 1, 9, 22, 9),
}
DIATOMIC = {
    # name: (Z1, Z2, Rbond, primbas, nnodes, nelem, lmax, mmax, lang, mang)
    "h2_sigma": (1, 1, 1.4, 4, 6, 2, 2, 0, 20, 5),
    "n2_l2m1": (7, 7, 2.068, 4, 5, 2, 2, 1, 20, 9),
    "co_l3m2": (6, 8, 2.132, 4, 5, 2, [3, 3, 2], 2, 24, 13),
}


def atomic_oracle(case):
    """Grid2DOracle of an atomic case from the numpy set-up chain (oracle/fem_oracle.py) + (lval, mval, nrad, rs)."""
    Z, primbas, nnodes, nelem, lmax, mmax, nocc, lang, mang = ATOMIC[case]
    rs = fo.RadialSetup(primbas, nnodes, fo.get_grid(40.0, nelem), taylor_order=-1)
    # angular ordering of atomic/basis.cpp:185-195: m = 0 first, then +m, -m per |m|
    lval, mval = [], []
    for l in range(lmax + 1):
        lval.append(l); mval.append(0)
    for m in range(1, mmax + 1):
        for l in range(m, lmax + 1):
            lval.append(l); mval.append(m)
            lval.append(l); mval.append(-m)
    r = [rs.r(e) for e in range(rs.nel)]
    w = [rs.wrad(e) for e in range(rs.nel)]
    bf = [rs.get_bf(e) for e in range(rs.nel)]
    df = [rs.get_df(e) for e in range(rs.nel)]
    o = oracle.Grid2DOracle(0, lval, mval, rs.first, rs.last, r, w, bf, df, lang, mang)
    return o, np.array(lval), np.array(mval), rs


def atomic_matrices(rs, lval, Z):
    """S, T, Vnuc of the atomic product basis from the radial matrices (TwoDBasis.cpp:371-470)."""
    nang, nrad = len(lval), rs.nbf
    S = np.kron(np.eye(nang), rs.overlap())
    T = np.zeros_like(S)
    for a, l in enumerate(lval):
        T[a * nrad:(a + 1) * nrad, a * nrad:(a + 1) * nrad] = rs.kinetic() + l * (l + 1) * rs.kinetic_l()
    V = np.kron(np.eye(nang), rs.nuclear(Z))
    return S, T, V


def atomic_density(case, spin=None):
    Z, primbas, nnodes, nelem, lmax, mmax, nocc, lang, mang = ATOMIC[case]
    o, lval, mval, rs = atomic_oracle(case)
    S, T, V = atomic_matrices(rs, lval, Z)
    E, C = scipy.linalg.eigh(T + V, S)
    no = nocc if spin != "b" else nocc - 1
    P = (1.0 if spin else 2.0) * C[:, :no] @ C[:, :no].T
    # multiplicative seeded symmetric perturbation + a small dense additive one that couples every angular block
    n = P.shape[0]
    R = splitmix64(n * n, seed=20261104 + (3 if spin == "b" else 0)).reshape(n, n)
    R = 0.5 * (R + R.T)
    P = P * (1.0 + 1e-3 * R)
    D = splitmix64(n * 3, seed=99 + (1 if spin == "b" else 0)).reshape(n, 3) * np.exp(-0.3 * (np.arange(n) % rs.nbf))[:, None]
    P = P + 1e-3 * (D @ D.T)
    return np.asfortranarray(P)


def diatomic_density(n, spin=None):
    """positive semi-definite low-rank density + small seeded symmetric noise populating every angular block"""
    C = splitmix64(n * 4, seed=7 + (5 if spin == "b" else 0)).reshape(n, 4) * np.exp(-0.05 * np.arange(n))[:, None]
    R = splitmix64(n * n, seed=20261104 + (3 if spin == "b" else 0)).reshape(n, n)
    return np.asfortranarray(C @ C.T + 1e-4 * 0.5 * (R + R.T))
