"""ORACLE package (test infrastructure).  CPU restatement of the reference's algorithm for the
Fock-build path.  May be imported ONLY by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs -- never by helfem_b200 (the product).

  fem_oracle.py        numpy restatement of the finite-element setup (tables + stored integrals)
  xc_oracle.hpp        XC functionals, complex-step derivatives
  hotpath_oracle.cpp   coulomb / exchange / eval_Fxc loops (liboracle.so, built by `make -C oracle`)
  chan_oracle.cpp      atomic / diatomic coulomb and exchange
  grid2d_oracle.cpp    atomic / diatomic DFTGrid::eval_Fxc on the angular x radial product grid

Parity pins: see DESIGN.md section "Oracle".  XC values are "parity unpinned" upstream (libxc is an
un-vendored, unversioned dependency of the reference and none of its tests holds an XC value).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None
c_double_p = ctypes.POINTER(ctypes.c_double)
c_int_p = ctypes.POINTER(ctypes.c_int)


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.oracle_gaunt.restype = ctypes.c_double
        _lib.oracle_gaunt.argtypes = [ctypes.c_int] * 6
    return _lib


def _p(a):
    return a.ctypes.data_as(c_double_p)


def _ip(a):
    return a.ctypes.data_as(c_int_p)


def _ptr_array(mats):
    arr = (c_double_p * len(mats))()
    for i, m in enumerate(mats):
        arr[i] = _p(m)
    return arr


def gaunt(L, M, l, m, lp, mp):
    return lib().oracle_gaunt(L, M, l, m, lp, mp)


def xc(func, rho, sigma=None, tau=None, thr=1e-12, polarized=False):
    """libxc-style arrays (spin fastest).  Returns exc, vrho, vsigma, vtau for ONE functional id."""
    rho = np.ascontiguousarray(rho, dtype=np.float64)
    n = rho.size // (2 if polarized else 1)
    exc = np.zeros(n)
    vrho = np.zeros_like(rho)
    vs = np.zeros(n * (3 if polarized else 1))
    vt = np.zeros_like(rho)
    sg = None if sigma is None else np.ascontiguousarray(sigma, dtype=np.float64)
    ta = None if tau is None else np.ascontiguousarray(tau, dtype=np.float64)
    fn = lib().oracle_xc_pol if polarized else lib().oracle_xc_unpol
    fn(int(func), int(n), ctypes.c_double(thr), _p(rho), _p(sg) if sg is not None else None,
       _p(ta) if ta is not None else None, _p(exc), _p(vrho), _p(vs), _p(vt))
    return exc, vrho, vs, vt


class SadatomOracle:
    """Holds the tables of one basis (from fem_oracle.RadialSetup or from the product's host library)
    and evaluates the three operators with the reference's algorithm on the CPU."""

    def __init__(self, first, last, r, wrad, bf, df, lmax):
        self.first = np.ascontiguousarray(first, dtype=np.int32)
        self.last = np.ascontiguousarray(last, dtype=np.int32)
        self.nel = len(self.first)
        self.nrad = int(self.last[-1]) + 1
        self.r = np.ascontiguousarray(np.concatenate(r))
        self.wrad = np.ascontiguousarray(np.concatenate(wrad))
        self.nq = len(r[0])
        self.bf = np.concatenate([np.asfortranarray(b).reshape(-1, order="F") for b in bf])
        self.df = np.concatenate([np.asfortranarray(b).reshape(-1, order="F") for b in df])
        self.lmax = lmax
        self.dL = self.dm1L = self.tei = self.ktei = None

    def set_tei(self, dL, dm1L, tei, ktei=None):
        """lists indexed [L*nel+iel] of Fortran-ordered arrays"""
        self.dL = [np.asfortranarray(a) for a in dL]
        self.dm1L = [np.asfortranarray(a) for a in dm1L]
        self.tei = [np.asfortranarray(a) for a in tei]
        self.ktei = None if ktei is None else [np.asfortranarray(a) for a in ktei]

    def coulomb(self, P):
        P = np.asfortranarray(P, dtype=np.float64)
        J = np.zeros_like(P, order="F")
        n = self.nel
        lib().oracle_sadatom_coulomb(n, _ip(self.first), _ip(self.last), _ptr_array(self.dL[:n]), _ptr_array(self.dm1L[:n]),
                                     _ptr_array(self.tei[:n]), _p(P), self.nrad, _p(J))
        return J

    def exchange(self, P):
        P = np.asfortranarray(P, dtype=np.float64)
        K = np.zeros_like(P, order="F")
        lib().oracle_sadatom_exchange(self.lmax, self.nel, _ip(self.first), _ip(self.last), _ptr_array(self.dL),
                                      _ptr_array(self.dm1L), _ptr_array(self.ktei), _p(P), self.nrad, _p(K))
        return K

    def eval_fxc(self, x_func, c_func, P, Pb=None, thr=1e-12, beta=True):
        P = np.asfortranarray(P, dtype=np.float64)
        H = np.zeros_like(P, order="F")
        exc, nel = ctypes.c_double(), ctypes.c_double()
        nspin = 1 if Pb is None else 2
        Hb = None
        if Pb is not None:
            Pb = np.asfortranarray(Pb, dtype=np.float64)
            Hb = np.zeros_like(P, order="F")
        rc = lib().oracle_sadatom_eval_fxc(self.nel, self.nq, _ip(self.first), _ip(self.last), _p(self.r), _p(self.wrad),
                                           _p(self.bf), _p(self.df), int(x_func), int(c_func), nspin, _p(P),
                                           _p(Pb) if Pb is not None else None, self.nrad, int(P.shape[2]),
                                           ctypes.c_double(thr), _p(H), _p(Hb) if Hb is not None else None,
                                           int(bool(beta)), ctypes.byref(exc), ctypes.byref(nel))
        if rc != 0:
            raise RuntimeError("oracle: functional not restated")
        if Pb is None:
            return H, exc.value, nel.value
        return H, Hb, exc.value, nel.value


def sadatom_from_setup(rs, lmax, nL=1, exchange=False):
    """Build a SadatomOracle (tables + stored integrals) from fem_oracle.RadialSetup."""
    from . import fem_oracle
    o = SadatomOracle(rs.first, rs.last, [rs.r(e) for e in range(rs.nel)], [rs.wrad(e) for e in range(rs.nel)],
                      [rs.get_bf(e) for e in range(rs.nel)], [rs.get_df(e) for e in range(rs.nel)], lmax)
    dL, dm1L, tei, ktei = [], [], [], []
    for L in range(nL):
        for e in range(rs.nel):
            dL.append(rs.radial_integral(L, e))
            dm1L.append(rs.radial_integral(-L - 1, e))
            t = rs.twoe_integral(L, e)
            tei.append(t)
            if exchange:
                n = rs.basis(e).nbf
                ktei.append(fem_oracle.exchange_tei(t, n, n, n, n))
    o.set_tei(dL, dm1L, tei, ktei if exchange else None)
    return o


class AtomicOracle:
    """Reference algorithm of atomic::basis::TwoDBasis::coulomb / exchange on the CPU (oracle/chan_oracle.cpp).
    dL, dm1L, tei, ktei: lists indexed [L*nel + iel] of Fortran-ordered arrays."""

    def __init__(self, lval, mval, first, last, dL, dm1L, tei, ktei=None):
        self.lval = np.ascontiguousarray(lval, dtype=np.int32)
        self.mval = np.ascontiguousarray(mval, dtype=np.int32)
        self.first = np.ascontiguousarray(first, dtype=np.int32)
        self.last = np.ascontiguousarray(last, dtype=np.int32)
        self.nel, self.nang = len(self.first), len(self.lval)
        self.dL = [np.asfortranarray(a) for a in dL]
        self.dm1L = [np.asfortranarray(a) for a in dm1L]
        self.tei = [np.asfortranarray(a) for a in tei]
        self.ktei = None if ktei is None else [np.asfortranarray(a) for a in ktei]

    def coulomb(self, P):
        P = np.asfortranarray(P, dtype=np.float64)
        J = np.zeros_like(P, order="F")
        lib().oracle_atomic_coulomb(self.nang, _ip(self.lval), _ip(self.mval), self.nel, _ip(self.first), _ip(self.last),
                                    _ptr_array(self.dL), _ptr_array(self.dm1L), _ptr_array(self.tei), _p(P), _p(J))
        return J

    def exchange(self, P):
        P = np.asfortranarray(P, dtype=np.float64)
        K = np.zeros_like(P, order="F")
        lib().oracle_atomic_exchange(self.nang, _ip(self.lval), _ip(self.mval), self.nel, _ip(self.first), _ip(self.last),
                                     _ptr_array(self.dL), _ptr_array(self.dm1L), _ptr_array(self.ktei), _p(P), _p(K))
        return K


class DiatomicOracle:
    """Reference algorithm of diatomic::basis::TwoDBasis::coulomb / exchange on the CPU (oracle/chan_oracle.cpp),
    in the dummy space; `pure` gathers / scatters as expand_boundaries / remove_boundaries do.
    dis: 4 lists (P0, P2, Q0, Q2) indexed [ilm*nel+iel]; tei, ktei: lists indexed [4*(ilm*nel+iel) + 2a+b]."""

    def __init__(self, lval, mval, first, last, Lmap, Mmap, Rhalf, dis, tei, ktei, pure):
        ci = lambda a: np.ascontiguousarray(a, dtype=np.int32)
        self.lval, self.mval, self.first, self.last, self.Lmap, self.Mmap = map(ci, (lval, mval, first, last, Lmap, Mmap))
        self.Rhalf = float(Rhalf)
        self.dis = [[np.asfortranarray(a) for a in d] for d in dis]
        self.tei = None if tei is None else [np.asfortranarray(a) for a in tei]
        self.ktei = None if ktei is None else [np.asfortranarray(a) for a in ktei]
        self.pure = np.asarray(pure)
        self.nd = len(self.lval) * (int(self.last[-1]) + 1)

    def _apply(self, fn, teis, Ppure):
        P = np.zeros((self.nd, self.nd), order="F")
        P[np.ix_(self.pure, self.pure)] = Ppure
        F = np.zeros_like(P, order="F")
        rc = fn(len(self.lval), _ip(self.lval), _ip(self.mval), len(self.first), _ip(self.first), _ip(self.last), len(self.Lmap),
                _ip(self.Lmap), _ip(self.Mmap), ctypes.c_double(self.Rhalf), _ptr_array(self.dis[0]), _ptr_array(self.dis[1]),
                _ptr_array(self.dis[2]), _ptr_array(self.dis[3]), _ptr_array(teis), _p(P), _p(F))
        assert rc == 0
        return np.asfortranarray(F[np.ix_(self.pure, self.pure)])

    def coulomb(self, P):
        return self._apply(lib().oracle_diatomic_coulomb, self.tei, P)

    def exchange(self, P):
        return self._apply(lib().oracle_diatomic_exchange, self.ktei, P)


    def _blocks(self, fn, teis, Ppure, alist, blist):
        """Block-restricted oracle (chan_oracle.cpp): the listed output angular blocks only, from the caller's
        PURE-basis matrix (no dense dummy-space copy): returns (nblk, nrad, nrad) blocks in the dummy space,
        block[q][a, b] = F[(alist[q], a), (blist[q], b)]."""
        P = np.asfortranarray(Ppure, dtype=np.float64)
        amap = np.full(self.nd, -1, dtype=np.int32)
        amap[self.pure] = np.arange(len(self.pure), dtype=np.int32)
        al, bl = np.ascontiguousarray(alist, dtype=np.int32), np.ascontiguousarray(blist, dtype=np.int32)
        nrad = int(self.last[-1]) + 1
        out = np.zeros((len(al), nrad * nrad))
        fn.argtypes = None
        rc = fn(len(self.lval), _ip(self.lval), _ip(self.mval), len(self.first), _ip(self.first), _ip(self.last), len(self.Lmap),
                _ip(self.Lmap), _ip(self.Mmap), ctypes.c_double(self.Rhalf), _ptr_array(self.dis[0]), _ptr_array(self.dis[1]),
                _ptr_array(self.dis[2]), _ptr_array(self.dis[3]), _ptr_array(teis), _p(P), ctypes.c_long(P.shape[0]), _ip(amap),
                len(al), _ip(al), _ip(bl), _p(out))
        assert rc == 0
        return out.reshape(len(al), nrad, nrad).transpose(0, 2, 1)  # column-major blocks -> [q][a, b]

    def exchange_blocks(self, P, jlist, klist):
        return self._blocks(lib().oracle_diatomic_exchange_blocks, self.ktei, P, jlist, klist)

    def coulomb_blocks(self, P, ilist, jlist):
        return self._blocks(lib().oracle_diatomic_coulomb_blocks, self.tei, P, ilist, jlist)

    def pure_block(self, F, a, b):
        """Block (a, b) of a PURE-basis matrix, expanded to the dummy space (zeros on the removed rows / columns)."""
        nrad = int(self.last[-1]) + 1
        amap = np.full(self.nd, -1, dtype=np.int64)
        amap[self.pure] = np.arange(len(self.pure))
        ra, rb = amap[a * nrad:(a + 1) * nrad], amap[b * nrad:(b + 1) * nrad]
        out = np.zeros((nrad, nrad))
        ia, ib = np.nonzero(ra >= 0)[0], np.nonzero(rb >= 0)[0]
        out[np.ix_(ia, ib)] = F[np.ix_(ra[ia], rb[ib])]
        return out


def angular_grid(lang, mang):
    """angular::angular_chebyshev (/root/reference/src/general/angular.cpp:22-45,64-71): Chebyshev rule in cos(theta)
    (fem_oracle.chebyshev) x uniform phi, point index i*nphi + j."""
    from . import fem_oracle
    x, w = fem_oracle.chebyshev(lang)
    dphi = 2.0 * np.pi / mang
    cth = np.repeat(x, mang)
    phi = np.tile(np.arange(mang) * dphi, lang)
    return cth, phi, np.repeat(w, mang) * dphi


def ylm_table(lval, mval, cth, phi, shift=0):
    """Y_l^{m+shift}(theta, phi) with the Condon-Shortley phase (what gsl_sf_legendre_sphPlm * e^{i m phi} gives,
    /root/reference/src/general/spherical_harmonics.cpp:25-41) from scipy; (nang x npts) complex, zero where m+shift > l."""
    from scipy.special import sph_harm_y
    th = np.arccos(cth)
    out = np.zeros((len(lval), len(cth)), dtype=np.complex128)
    for a, (l, m) in enumerate(zip(lval, mval)):
        if abs(m + shift) <= l:
            out[a] = sph_harm_y(int(l), int(m + shift), th, phi)
    return out


class Grid2DOracle:
    """Reference algorithm of atomic / diatomic DFTGrid::eval_Fxc on the CPU (oracle/grid2d_oracle.cpp), in the
    full product ("dummy") space.  kind: 0 atomic, 1 diatomic.  r, wrad, bf, df: per-element lists
    (radial.get_r / get_wrad / get_bf / get_df); pure: optional index list (expand / remove_boundaries)."""

    def __init__(self, kind, lval, mval, first, last, r, wrad, bf, df, lang, mang, Rhalf=0.0, pure=None):
        ci = lambda a: np.ascontiguousarray(a, dtype=np.int32)
        self.kind, self.Rhalf = int(kind), float(Rhalf)
        self.lval, self.mval, self.first, self.last = map(ci, (lval, mval, first, last))
        self.nel, self.nang = len(self.first), len(self.lval)
        self.nrad = int(self.last[-1]) + 1
        self.nq = len(r[0])
        self.r = np.ascontiguousarray(np.concatenate(r))
        self.wrad = np.ascontiguousarray(np.concatenate(wrad))
        self.bf = np.concatenate([np.asfortranarray(b).reshape(-1, order="F") for b in bf])
        self.df = np.concatenate([np.asfortranarray(b).reshape(-1, order="F") for b in df])
        self.cth, self.phi, self.wang = (np.ascontiguousarray(a) for a in angular_grid(lang, mang))
        # memory order [a + nang*ia] as (re, im) pairs
        self.Y = np.ascontiguousarray(ylm_table(self.lval, self.mval, self.cth, self.phi, 0).T)
        self.Y1 = np.ascontiguousarray(ylm_table(self.lval, self.mval, self.cth, self.phi, 1).T)
        self.pure = None if pure is None else np.asarray(pure)
        self.nd = self.nang * self.nrad

    def _expand(self, P):
        if self.pure is None:
            return np.asfortranarray(P, dtype=np.float64)
        out = np.zeros((self.nd, self.nd), order="F")
        out[np.ix_(self.pure, self.pure)] = P
        return out

    def _remove(self, H):
        return H if self.pure is None else np.asfortranarray(H[np.ix_(self.pure, self.pure)])

    def eval_fxc(self, x_func, c_func, P, Pb=None, thr=1e-12, beta=True):
        """returns (H, Exc, Nel, Ekin) or (Ha, Hb, Exc, Nel, Ekin)"""
        Pa = self._expand(P)
        nspin = 1 if Pb is None else 2
        Pbx = None if Pb is None else self._expand(Pb)
        Ha = np.zeros_like(Pa, order="F")
        Hb = None if Pb is None else np.zeros_like(Pa, order="F")
        exc, nel, ekin = ctypes.c_double(), ctypes.c_double(), ctypes.c_double()
        Yv = self.Y.view(np.float64)
        Y1v = self.Y1.view(np.float64)
        rc = lib().oracle_grid2d_eval_fxc(self.kind, ctypes.c_double(self.Rhalf), self.nang, _ip(self.lval), _ip(self.mval),
                                          self.nel, self.nq, _ip(self.first), _ip(self.last), self.nrad, _p(self.r),
                                          _p(self.wrad), _p(self.bf), _p(self.df), len(self.cth), _p(self.cth), _p(self.phi),
                                          _p(self.wang), _p(Yv), _p(Y1v), int(x_func), int(c_func), nspin, _p(Pa),
                                          _p(Pbx) if Pbx is not None else None, ctypes.c_double(thr), _p(Ha),
                                          _p(Hb) if Hb is not None else None, int(bool(beta)), ctypes.byref(exc),
                                          ctypes.byref(nel), ctypes.byref(ekin))
        if rc != 0:
            raise RuntimeError("oracle: functional not restated")
        if Pb is None:
            return self._remove(Ha), exc.value, nel.value, ekin.value
        return self._remove(Ha), self._remove(Hb), exc.value, nel.value, ekin.value
