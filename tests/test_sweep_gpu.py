"""GPU parity of the Z = 1..86 sadatom sweep (BASELINE configs[4], run with -m gpu): ONE shared radial basis (20 LIP-15 elements,
lmax 3; the basis does not depend on Z), every atom's Aufbau ground configuration as spin densities, one batched call through
the C ABI (hfb200_sadatom_fock_host: 86 unrestricted PBE Fock builds), compared atom by atom with the CPU oracle evaluated
on tables from the independent numpy setup (oracle/fem_oracle.py).  Reference path per atom: one gensap process,
/root/reference/src/sadatom/main.cpp:101-143, solver.cpp:758-849 (FockBuild: coulomb(P/4pi), eval_Fxc(Pl/4pi))."""
import numpy as np
import pytest

import oracle
from oracle import fem_oracle as fo
from util import sadatom_density

pytestmark = pytest.mark.gpu

TOL_F, TOL_E = 1e-9, 1e-10


def test_sweep_every_atom_matches_oracle(native):
    import helfem_b200 as hb
    primbas, nnodes, nelem, lmax = 4, 15, 20, 3
    b = hb.SadatomBasis(Z=1, nelem=nelem, nnodes=nnodes, primbas=primbas, taylor_order=-1, lmax=lmax)
    b.compute_tei(1, False)
    rs = fo.RadialSetup(primbas, nnodes, fo.get_grid(40.0, nelem), taylor_order=-1)
    o = oracle.sadatom_from_setup(rs, lmax, nL=1, exchange=False)
    S, T, Tl = rs.overlap(), rs.kinetic(), rs.kinetic_l()
    nsl, blk = lmax + 1, b.nel * b.np_max**2
    Zs = list(range(1, 87))
    dens = {}
    hP = np.zeros((len(Zs), 2, nsl, blk))
    for q, Z in enumerate(Zs):
        V = rs.nuclear(Z)
        Pa = sadatom_density(S, T, Tl, V, Z, lmax, perturb=0, spin="a") / (4 * np.pi)
        Pb = sadatom_density(S, T, Tl, V, Z, lmax, perturb=0, spin="b") / (4 * np.pi)
        dens[Z] = (Pa, Pb)
        for s, P in enumerate((Pa, Pb)):
            for l in range(nsl):
                hP[q, s, l] = b.pack_band(P[:, :, l])
    hJ = np.zeros((len(Zs), blk))
    hH = np.zeros((len(Zs), 2, nsl, blk))
    hS = np.zeros((len(Zs), 4))
    b.fock_host(len(Zs), 101, 130, 2, nsl, 1e-12, hP, hJ, hH, hS)
    worst_j = worst_h = worst_e = worst_n = 0.0
    for q, Z in enumerate(Zs):
        Pa, Pb = dens[Z]
        Jo = o.coulomb(np.asfortranarray((Pa + Pb).sum(axis=2)))
        Ha, Hb, exc, nel = o.eval_fxc(101, 130, Pa, Pb)
        tol = max(TOL_F, 5e-14 * float(max(np.max(np.abs(Ha)), np.max(np.abs(Hb)))))
        dj = float(np.max(np.abs(b.unpack_band(hJ[q]) - Jo)))
        dh = 0.0
        for s, Ho in enumerate((Ha, Hb)):
            for l in range(nsl):
                dh = max(dh, float(np.max(np.abs(b.unpack_band(hH[q, s, l]) - Ho[:, :, l]))))
        assert dj <= TOL_F, (Z, dj)
        assert dh <= tol, (Z, dh, tol)
        assert abs(hS[q, 0] - exc) <= TOL_E * max(1.0, abs(exc)), (Z, hS[q, 0], exc)
        assert abs(hS[q, 1] - nel) <= TOL_E * max(1.0, abs(nel)) and abs(nel - Z) <= 1e-8 * Z, (Z, hS[q, 1], nel)
        worst_j, worst_h = max(worst_j, dj), max(worst_h, dh)
        worst_e, worst_n = max(worst_e, abs(hS[q, 0] - exc)), max(worst_n, abs(hS[q, 1] - Z))
    print("sweep Z=1..86: max|dJ| = %.2e, max|dVxc| = %.2e, max|dExc| = %.2e, max|Nel - Z| = %.2e" % (worst_j, worst_h, worst_e, worst_n))
