"""SURVEY section 8 "config 4b" AT FULL SIZE on the GPU (run with -m gpu): diatomic::dftgrid::DFTGrid::eval_Fxc
(/root/reference/src/diatomic/dftgrid.cpp:769-856, compute_bf :669-755) for N2, lmax 24, mmax 6, 5 LIP-15 elements, PBE, on
the default product grid ldft = 4 lmax + 12 = 108, mdft = 4 (mmax + 1) + 5 = 33 (diatomic/main.cpp:417,425): 3 564 angular x 375 mu
points = 1.34 M points, 19 552 basis functions.

The reference-shaped oracle (oracle/grid2d_oracle.cpp: dense complex basis matrices, Pv = P conj(bf)) needs ~2e14 flop for the
density alone at this size, so the check uses the ORBITAL form of the same sums, in numpy: for P = 2 C C^T
    rho(p) = 2 sum_o |psi_o(p)|^2,   d_c rho = 4 Re sum_o conj(psi_o) d_c psi_o,    psi_o = sum_i C_io bf_i,
which is algebraically identical to the reference's rho = Re sum_ij P_ij conj(bf_j) bf_i and shares nothing with the product's
pair-table factorisation.  XC at the points comes from the oracle's functionals; Exc, Nel and 24 sampled Fock blocks (a, b)
(m = 0, +-1, +-6, lowest / highest l, equal and different m) are then summed directly with the reference's increment formulas
(dftgrid.h:190-250).  Spherical harmonics from scipy.  Gate: 1e-9 on Fock elements, 1e-10 relative on Exc and Nel."""
import gc

import numpy as np
import pytest

import oracle
from util import diatomic_orbitals

pytestmark = pytest.mark.gpu

CFG = dict(Z1=7, Z2=7, Rbond=2.068, nelem=5, nnodes=15, primbas=4, lmax=24, mmax=6)


def numpy_reference(b, C, lang, mang, blocks):
    nang, nrad = b.nang, b.nrad
    lval, mval = np.asarray(b.lval), np.asarray(b.mval)
    cth, phi, wang = oracle.angular_grid(lang, mang)
    npts = len(cth)
    Y = oracle.ylm_table(lval, mval, cth, phi, 0)      # (nang, npts)
    Y1 = oracle.ylm_table(lval, mval, cth, phi, 1)
    sth = np.sqrt(1.0 - cth**2)
    # d/dtheta Y_lm = m cot(theta) Y_lm + sqrt((l-m)(l+m+1)) e^{-i phi} Y_l^{m+1}   (TwoDBasis.cpp:1407-1419)
    dY = mval[:, None] * (cth / sth)[None, :] * Y + np.sqrt(((lval - mval) * (lval + mval + 1)).astype(float))[:, None] * np.exp(-1j * phi)[None, :] * Y1
    pY = 1j * mval[:, None] * Y
    Cc = C.reshape(nang, nrad, -1)
    norb = Cc.shape[2]
    Rh = b.Rhalf
    exc_sum = nel_sum = 0.0
    Hb = {ab: np.zeros((nrad, nrad)) for ab in blocks}
    for e in range(b.nel):
        mu, wr, bf, df = b.radial(e)
        f0, l0 = int(b.first[e]), int(b.last[e])
        nq = len(mu)
        Ce = Cc[:, f0:l0 + 1, :]
        G0 = np.einsum("qk,ako->aqo", bf, Ce).reshape(nang, nq * norb)
        G1 = np.einsum("qk,ako->aqo", df, Ce).reshape(nang, nq * norb)
        psi = (Y.T @ G0).reshape(npts, nq, norb)
        psi_m = (Y.T @ G1).reshape(npts, nq, norb)
        psi_n = (dY.T @ G0).reshape(npts, nq, norb)
        psi_p = (pY.T @ G0).reshape(npts, nq, norb)
        sh = np.sinh(mu)
        hm = Rh * np.sqrt(sh[None, :] ** 2 + sth[:, None] ** 2)   # h_mu = h_nu
        hp = Rh * sh[None, :] * sth[:, None]
        w = wang[:, None] * wr[None, :] * Rh**3 * sh[None, :] * (sh[None, :] ** 2 + sth[:, None] ** 2)
        rho = 2.0 * np.sum(np.abs(psi) ** 2, axis=2)
        gm = 4.0 * np.sum((np.conj(psi) * psi_m).real, axis=2) / hm
        gn = 4.0 * np.sum((np.conj(psi) * psi_n).real, axis=2) / hm
        gp = 4.0 * np.sum((np.conj(psi) * psi_p).real, axis=2) / hp
        sigma = gm**2 + gn**2 + gp**2
        del psi, psi_m, psi_n, psi_p
        ex = np.zeros(rho.size)
        vr = np.zeros(rho.size)
        vs = np.zeros(rho.size)
        for fid in (101, 130):
            e_, vr_, vs_, _ = oracle.xc(fid, rho.ravel(), sigma.ravel(), np.zeros(rho.size), thr=1e-12, polarized=False)
            ex += e_; vr += vr_; vs += vs_
        ex, vr, vs = ex.reshape(rho.shape), vr.reshape(rho.shape), vs.reshape(rho.shape)
        exc_sum += float(np.sum(w * ex * rho))
        nel_sum += float(np.sum(w * rho))
        A = w * vr
        g_m, g_n, g_p = 2 * w * vs * gm / hm, 2 * w * vs * gn / hm, 2 * w * vs * gp / hp
        for (a, c) in blocks:
            yy = Y[a][:, None] * np.conj(Y[c])[:, None]                    # (npts, 1)
            ty = dY[a][:, None] * np.conj(Y[c])[:, None] + Y[a][:, None] * np.conj(dY[c])[:, None]
            V = np.sum((A * yy + g_n * ty + g_p * (1j * (mval[a] - mval[c])) * yy).real, axis=0)   # (nq,)
            Gq = np.sum((g_m * yy).real, axis=0)
            blk = bf.T @ (V[:, None] * bf) + df.T @ (Gq[:, None] * bf) + bf.T @ (Gq[:, None] * df)
            Hb[(a, c)][f0:l0 + 1, f0:l0 + 1] += blk
    return exc_sum, nel_sum, Hb


def test_cfg4b_eval_fxc_vs_orbital_reference(native):
    import helfem_b200 as hb
    b = hb.DiatomicBasis(CFG["Z1"], CFG["Z2"], CFG["Rbond"], nelem=CFG["nelem"], nnodes=CFG["nnodes"], primbas=CFG["primbas"],
                         lmax=CFG["lmax"], mmax=CFG["mmax"])
    lang, mang = 4 * CFG["lmax"] + 12, 4 * (CFG["mmax"] + 1) + 5
    assert (b.nang, b.nbf, lang * mang) == (283, 19552, 3564)
    g = hb.DiatomicDFTGrid(b, lang, mang)
    C = diatomic_orbitals(b)
    pure = b.pure_indices()
    P = np.asfortranarray(2.0 * C[pure] @ C[pure].T)
    H, exc, nel, ekin = g.eval_Fxc(101, 130, P)
    by_m = {m: [a for a in range(b.nang) if b.mval[a] == m] for m in range(-6, 7)}
    lo = {m: min(v, key=lambda a: b.lval[a]) for m, v in by_m.items()}
    hi = {m: max(v, key=lambda a: b.lval[a]) for m, v in by_m.items()}
    blocks = []
    for m in (0, 1, -1, 6, -6):
        blocks += [(lo[m], lo[m]), (lo[m], hi[m]), (hi[m], hi[m]), (by_m[m][len(by_m[m]) // 2], lo[m])]
    blocks += [(lo[0], lo[1]), (lo[1], lo[-1]), (hi[0], hi[6]), (by_m[0][3], by_m[2][2])]   # m_a != m_b: vanish for a linear molecule
    exc_o, nel_o, Hb = numpy_reference(b, C, lang, mang, blocks)
    amap = np.full(b.ndummy, -1, dtype=np.int64)
    amap[pure] = np.arange(len(pure))
    worst = scale = 0.0
    for (a, c) in blocks:
        ra, rc = amap[a * b.nrad:(a + 1) * b.nrad], amap[c * b.nrad:(c + 1) * b.nrad]
        ia, ic = np.nonzero(ra >= 0)[0], np.nonzero(rc >= 0)[0]
        got = H[np.ix_(ra[ia], rc[ic])]
        ref = Hb[(a, c)][np.ix_(ia, ic)]
        worst = max(worst, float(np.max(np.abs(got - ref))))
        scale = max(scale, float(np.max(np.abs(ref))))
        if b.mval[a] != b.mval[c]:
            assert float(np.max(np.abs(got))) <= 1e-10
    print("cfg4b eval_Fxc: Exc = %.10f (ref %.10f), Nel = %.10f (ref %.10f), %d blocks max|dVxc| = %.3e, max|Vxc| = %.3e"
          % (exc, exc_o, nel, nel_o, len(blocks), worst, scale))
    assert abs(nel - nel_o) <= 1e-10 * abs(nel_o) and abs(nel_o - 14.0) <= 1e-6
    assert abs(exc - exc_o) <= 1e-10 * abs(exc_o)
    assert scale > 1e-2 and worst <= 1e-9
    del H, P
    gc.collect()
