"""ctypes view of hfb::sadatom::TwoDBasis / DFTGrid (helfem_b200/host/sadatom.h), the host-side
mirror of helfem::sadatom::basis::TwoDBasis and helfem::sadatom::dftgrid::DFTGrid
(/root/reference/src/sadatom/basis.h:38-170, dftgrid.h:150-176).  Harness only."""
import ctypes

import numpy as np

from . import lib, check, check_host, _p, _ip


class SadatomBasis:
    """One spherically averaged atom basis; arguments follow the reference's `gensap` flags
    (/root/reference/src/sadatom/main.cpp:101-143)."""

    def __init__(self, Z, nelem, nnodes=15, primbas=4, nquad=0, rmax=40.0, igrid=4, zexp=2.0, taylor_order=-1,
                 lmax=3, zeroder=False, device=0):
        L = lib()
        self._h = ctypes.c_void_p()
        check_host(L.hfh_sadatom_new(ctypes.byref(self._h), int(Z), int(primbas), int(nnodes), int(nquad), int(nelem),
                                     ctypes.c_double(rmax), int(igrid), ctypes.c_double(zexp), int(taylor_order),
                                     int(lmax), int(bool(zeroder)), int(device)))
        nbf, nel, nq, npm = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        tc, td = ctypes.c_double(), ctypes.c_double()
        check_host(L.hfh_sadatom_dims(self._h, ctypes.byref(nbf), ctypes.byref(nel), ctypes.byref(nq), ctypes.byref(npm),
                                      ctypes.byref(tc), ctypes.byref(td)))
        self.Z, self.lmax = Z, lmax
        self.nbf, self.nel, self.nq, self.np_max = nbf.value, nel.value, nq.value, npm.value
        self.taylor_cutoff, self.taylor_diff = tc.value, td.value
        self.first = np.zeros(self.nel, dtype=np.int32)
        self.last = np.zeros(self.nel, dtype=np.int32)
        check_host(L.hfh_sadatom_idx(self._h, _ip(self.first), _ip(self.last)))
        self.nL = 0

    def __del__(self):
        try:
            if self._h:
                lib().hfh_sadatom_free(self._h)
                self._h = None
        except Exception:
            pass

    # ---- one-electron matrices (setup side) ----
    def _matrix(self, which):
        out = np.zeros((self.nbf, self.nbf), order="F")
        check_host(lib().hfh_sadatom_matrix(self._h, which, _p(out)))
        return out

    def overlap(self):
        return self._matrix(0)

    def kinetic(self):
        return self._matrix(1)

    def kinetic_l(self):
        return self._matrix(2)

    def nuclear(self):
        return self._matrix(3)

    def tables(self, iel):
        """r, wrad (nq) and bf, df, lf (nq x np_iel, Fortran order) of one element."""
        npe = int(self.last[iel] - self.first[iel] + 1)
        r = np.zeros(self.nq)
        w = np.zeros(self.nq)
        bf = np.zeros((self.nq, npe), order="F")
        df = np.zeros((self.nq, npe), order="F")
        lf = np.zeros((self.nq, npe), order="F")
        check_host(lib().hfh_sadatom_tables(self._h, int(iel), _p(r), _p(w), _p(bf), _p(df), _p(lf)))
        return r, w, bf, df, lf

    def compute_tei(self, nL=0, exchange=True):
        check_host(lib().hfh_sadatom_compute_tei(self._h, int(nL), int(bool(exchange))))
        self.nL = nL if nL > 0 else 2 * self.lmax + 1
        self.have_ktei = bool(exchange)

    def tei(self, L, iel):
        """disjoint_L, disjoint_m1L (np x np), prim_tei [, prim_ktei] (np^2 x np^2)."""
        npe = int(self.last[iel] - self.first[iel] + 1)
        a = np.zeros((npe, npe), order="F")
        b = np.zeros((npe, npe), order="F")
        t = np.zeros((npe * npe, npe * npe), order="F")
        k = np.zeros((npe * npe, npe * npe), order="F") if self.have_ktei else None
        check_host(lib().hfh_sadatom_get_tei(self._h, int(L), int(iel), _p(a), _p(b), _p(t), _p(k) if k is not None else None))
        return a, b, t, k

    # ---- the hot path (GPU through the C ABI) ----
    def context(self):
        ctx = ctypes.c_void_p()
        check_host(lib().hfh_sadatom_context(self._h, ctypes.byref(ctx)))
        return ctx

    def coulomb(self, P):
        P = np.asfortranarray(P, dtype=np.float64)
        J = np.zeros_like(P, order="F")
        check_host(lib().hfh_sadatom_coulomb(self._h, _p(P), _p(J)))
        return J

    def exchange(self, P):
        """P: (nbf, nbf, lmax+1) Fortran-ordered cube; returns -K like the reference."""
        P = np.asfortranarray(P, dtype=np.float64)
        K = np.zeros_like(P, order="F")
        check_host(lib().hfh_sadatom_exchange(self._h, _p(P), int(P.shape[2]), _p(K)))
        return K

    def eval_fxc(self, x_func, c_func, P, Pb=None, thr=1e-12, beta=True):
        """Restricted (Pb None) or unrestricted eval_Fxc.  Returns (H[, Hb], Exc, Nel)."""
        P = np.asfortranarray(P, dtype=np.float64)
        exc, nel = ctypes.c_double(), ctypes.c_double()
        H = np.zeros_like(P, order="F")
        if Pb is None:
            check_host(lib().hfh_sadatom_eval_fxc(self._h, int(x_func), int(c_func), 1, _p(P), None, int(P.shape[2]),
                                                  ctypes.c_double(thr), _p(H), None, 0, ctypes.byref(exc), ctypes.byref(nel)))
            return H, exc.value, nel.value
        Pb = np.asfortranarray(Pb, dtype=np.float64)
        Hb = np.zeros_like(P, order="F")
        check_host(lib().hfh_sadatom_eval_fxc(self._h, int(x_func), int(c_func), 2, _p(P), _p(Pb), int(P.shape[2]),
                                              ctypes.c_double(thr), _p(H), _p(Hb), int(bool(beta)), ctypes.byref(exc),
                                              ctypes.byref(nel)))
        return H, Hb, exc.value, nel.value

    # ---- element-band helpers for the batched device API ----
    def pack_band(self, M):
        M = np.asfortranarray(M, dtype=np.float64)
        band = np.zeros(self.nel * self.np_max * self.np_max)
        check(lib().hfb200_sadatom_pack_band(self.context(), _p(M), int(M.shape[0]), _p(band)))
        return band

    def unpack_band(self, band):
        M = np.zeros((self.nbf, self.nbf), order="F")
        band = np.ascontiguousarray(band, dtype=np.float64)
        check(lib().hfb200_sadatom_unpack_band_add(self.context(), _p(band), _p(M), int(self.nbf)))
        return M

    def fock_batch(self, nbatch, x_func, c_func, nspin, nslices, thr, dP, dJ, dH, dS, stream=0):
        """Device pointers (ints) of element-band buffers; see include/helfem_b200.h."""
        check(lib().hfb200_sadatom_fock_batch(self.context(), int(nbatch), int(x_func), int(c_func), int(nspin), int(nslices),
                                              ctypes.c_double(thr), ctypes.c_void_p(dP), ctypes.c_void_p(dJ), ctypes.c_void_p(dH),
                                              ctypes.c_void_p(dS), ctypes.c_void_p(stream)))

    def fock_batch_static(self, nbatch, x_func, c_func, nspin, nslices, thr, dP, dJ, dH, dS, stream=0):
        """fock_batch without any allocation inside the call (static device scratch, bounded batch size)."""
        check(lib().hfb200_sadatom_fock_batch_static(self.context(), int(nbatch), int(x_func), int(c_func), int(nspin),
                                                     int(nslices), ctypes.c_double(thr), ctypes.c_void_p(dP), ctypes.c_void_p(dJ),
                                                     ctypes.c_void_p(dH), ctypes.c_void_p(dS), ctypes.c_void_p(stream)))

    def fock_host(self, nbatch, x_func, c_func, nspin, nslices, thr, hP, hJ=None, hH=None, hS=None):
        """Host element-band arrays (numpy, C-contiguous float64) in and out; see include/helfem_b200.h."""
        check(lib().hfb200_sadatom_fock_host(self.context(), int(nbatch), int(x_func), int(c_func), int(nspin), int(nslices),
                                             ctypes.c_double(thr), _p(hP), _p(hJ) if hJ is not None else None,
                                             _p(hH) if hH is not None else None, _p(hS) if hS is not None else None))
