"""ctypes view of hfb::atomic::TwoDBasis (helfem_b200/host/atomic.h), the host-side mirror of
helfem::atomic::basis::TwoDBasis (/root/reference/src/atomic/TwoDBasis.h:38-226).  Harness only."""
import ctypes

import numpy as np

from . import lib, check, check_host, _p, _ip


class AtomicBasis:
    """Arguments follow the reference's `atomic` flags (/root/reference/src/atomic/main.cpp:63-119)."""

    def __init__(self, Z, nelem, nnodes=15, primbas=4, nquad=0, rmax=40.0, igrid=4, zexp=2.0, taylor_order=-1,
                 lmax=0, mmax=0, zeroder=False, device=0):
        L = lib()
        L.hfh_atomic_free.argtypes = [ctypes.c_void_p]
        L.hfh_atomic_free.restype = None
        self._h = ctypes.c_void_p()
        check_host(L.hfh_atomic_new(ctypes.byref(self._h), int(Z), int(primbas), int(nnodes), int(nquad), int(nelem),
                                    ctypes.c_double(rmax), int(igrid), ctypes.c_double(zexp), int(taylor_order),
                                    int(lmax), int(mmax), int(bool(zeroder)), int(device)))
        nbf, nrad, nang, nel, nL = (ctypes.c_int() for _ in range(5))
        check_host(L.hfh_atomic_dims(self._h, ctypes.byref(nbf), ctypes.byref(nrad), ctypes.byref(nang), ctypes.byref(nel),
                                     ctypes.byref(nL)))
        self.Z = Z
        self.nbf, self.nrad, self.nang, self.nel = nbf.value, nrad.value, nang.value, nel.value
        self.lval = np.zeros(self.nang, dtype=np.int32)
        self.mval = np.zeros(self.nang, dtype=np.int32)
        check_host(L.hfh_atomic_lm(self._h, _ip(self.lval), _ip(self.mval)))
        self.first = np.zeros(self.nel, dtype=np.int32)
        self.last = np.zeros(self.nel, dtype=np.int32)
        check_host(L.hfh_atomic_idx(self._h, _ip(self.first), _ip(self.last)))
        self.nL = 0
        self.have_ktei = False

    def __del__(self):
        try:
            if self._h:
                lib().hfh_atomic_free(self._h)
                self._h = None
        except Exception:
            pass

    def _matrix(self, which):
        out = np.zeros((self.nbf, self.nbf), order="F")
        check_host(lib().hfh_atomic_matrix(self._h, which, _p(out)))
        return out

    def overlap(self):
        return self._matrix(0)

    def kinetic(self):
        return self._matrix(1)

    def nuclear(self):
        return self._matrix(3)

    def compute_tei(self, exchange=True):
        check_host(lib().hfh_atomic_compute_tei(self._h, int(bool(exchange))))
        self.nL = 2 * int(self.lval.max()) + 1
        self.have_ktei = bool(exchange)

    def tei(self, L, iel):
        npe = int(self.last[iel] - self.first[iel] + 1)
        a = np.zeros((npe, npe), order="F")
        b = np.zeros((npe, npe), order="F")
        t = np.zeros((npe * npe, npe * npe), order="F")
        k = np.zeros((npe * npe, npe * npe), order="F") if self.have_ktei else None
        check_host(lib().hfh_atomic_get_tei(self._h, int(L), int(iel), _p(a), _p(b), _p(t), _p(k) if k is not None else None))
        return a, b, t, k

    def context(self):
        """The hfb200_ctx of this basis (device tables are uploaded on first use)."""
        ctx = ctypes.c_void_p()
        check_host(lib().hfh_atomic_context(self._h, ctypes.byref(ctx)))
        return ctx

    def chan_dim(self):
        n = ctypes.c_int()
        check(lib().hfb200_chan_dim(self.context(), ctypes.byref(n)))
        return n.value

    def coulomb_device(self, dP, dJ, n, stream=0):
        """Device-resident coulomb: dP, dJ = device pointers of dense (n x n) matrices, n = chan_dim()."""
        check(lib().hfb200_chan_coulomb_device(self.context(), ctypes.c_void_p(dP), int(n), ctypes.c_void_p(dJ), int(n),
                                                ctypes.c_void_p(stream)))

    def exchange_device(self, dP, dK, n, stream=0):
        check(lib().hfb200_chan_exchange_device(self.context(), ctypes.c_void_p(dP), int(n), ctypes.c_void_p(dK), int(n),
                                                 ctypes.c_void_p(stream)))

    def fock_device(self, dP, dJ, dK, n, stream=0):
        """J and K of one Fock build from the same device-resident P (Coulomb kernels on a second stream next to the exchange pipeline)."""
        check(lib().hfb200_chan_fock_device(self.context(), ctypes.c_void_p(dP), int(n), ctypes.c_void_p(dJ), int(n), ctypes.c_void_p(dK), int(n),
                                             ctypes.c_void_p(stream)))

    def exchange_stats(self):
        items, fl, ms = ctypes.c_longlong(), ctypes.c_double(), (ctypes.c_double * 5)()
        check(lib().hfb200_chan_exchange_stats(self.context(), ctypes.byref(items), ctypes.byref(fl), ms))
        return items.value, fl.value, list(ms)

    def exchange_batches(self):
        n = ctypes.c_int()
        check(lib().hfb200_chan_exchange_batches(self.context(), ctypes.byref(n)))
        return n.value

    def set_profile(self, on=True):
        check(lib().hfb200_chan_set_profile(self.context(), int(bool(on))))

    def set_shard(self, rank, nranks):
        """Multi-GPU: contract only the integral families r % nranks == rank (before the first operator call)."""
        check_host(lib().hfh_atomic_set_shard(self._h, int(rank), int(nranks)))

    def exchange_counts(self):
        out = (ctypes.c_longlong * 12)()
        check(lib().hfb200_chan_exchange_counts(self.context(), out, 12))
        return list(out)

    def allreduce_ms(self):
        ms = ctypes.c_double()
        check(lib().hfb200_chan_allreduce_ms(self.context(), ctypes.byref(ms)))
        return ms.value

    def comm_init(self, uid, rank, nranks):
        """Attach an NCCL communicator (id from helfem_b200.comm_unique_id() of rank 0) to the sharded context: the operators
        then return the complete J / K on every rank (all-reduce inside the library)."""
        check(lib().hfb200_comm_init(self.context(), ctypes.c_char_p(uid), int(rank), int(nranks)))

    def comm_set_root(self, root):
        """Root-driven host-matrix calls: after this (on every rank) only rank `root` calls coulomb / exchange; the others call
        serve(exchange) once per root call.  root < 0: back to the default (every rank passes the matrices)."""
        check(lib().hfb200_comm_set_root(self.context(), int(root)))

    def serve(self, exchange):
        """worker side of one root-driven coulomb (False) / exchange (True) call"""
        check(lib().hfb200_chan_serve(self.context(), int(bool(exchange))))

    def coulomb_into(self, P, J):
        """coulomb with a caller-owned output (Fortran-ordered float64, fully overwritten): no allocation per call."""
        check_host(lib().hfh_atomic_coulomb(self._h, _p(P), _p(J)))

    def exchange_into(self, P, K):
        check_host(lib().hfh_atomic_exchange(self._h, _p(P), _p(K)))

    def coulomb(self, P):
        P = np.asfortranarray(P, dtype=np.float64)
        J = np.zeros_like(P, order="F")
        check_host(lib().hfh_atomic_coulomb(self._h, _p(P), _p(J)))
        return J

    def exchange(self, P):
        """returns -K like the reference"""
        P = np.asfortranarray(P, dtype=np.float64)
        K = np.zeros_like(P, order="F")
        check_host(lib().hfh_atomic_exchange(self._h, _p(P), _p(K)))
        return K


class _ProductGrid:
    """ctypes view of hfb::{atomic,diatomic}::DFTGrid (helfem_b200/host/dftgrid.h), mirrors of
    helfem::{atomic,diatomic}::dftgrid::DFTGrid (/root/reference/src/atomic/dftgrid.h:160-187)."""
    _prefix = None

    def __init__(self, basis, lang, mang):
        L = lib()
        self._basis = basis
        self._h = ctypes.c_void_p()
        getattr(L, self._prefix + "_grid_free").argtypes = [ctypes.c_void_p]
        getattr(L, self._prefix + "_grid_free").restype = None
        check_host(getattr(L, self._prefix + "_grid_new")(ctypes.byref(self._h), basis._h, int(lang), int(mang)))
        self.lang, self.mang = lang, mang

    def __del__(self):
        try:
            if self._h:
                getattr(lib(), self._prefix + "_grid_free")(self._h)
                self._h = None
        except Exception:
            pass

    def eval_Fxc(self, x_func, c_func, P, Pb=None, thr=1e-12, beta=True, out=None):
        """returns (H, Exc, Nel, Ekin) or (Ha, Hb, Exc, Nel, Ekin); out: optional caller-owned Fortran-ordered float64 H (restricted
        call), fully overwritten -- no allocation per call"""
        P = np.asfortranarray(P, dtype=np.float64)
        if P.ndim != 2 or P.shape[0] != P.shape[1] or (Pb is not None and np.shape(Pb) != P.shape):
            raise ValueError("density matrices must be square and of equal shape")
        if out is not None and (Pb is not None or out.shape != P.shape or not out.flags.f_contiguous or out.dtype != np.float64):
            raise ValueError("out must be a Fortran-ordered float64 matrix of the shape of P (restricted call)")
        H = out if out is not None else np.empty_like(P, order="F")
        sc = np.zeros(3)
        fn = getattr(lib(), self._prefix + "_grid_eval")
        if Pb is None:
            check_host(fn(self._h, self._basis._h, int(P.shape[0]), int(x_func), int(c_func), 1, _p(P), None, ctypes.c_double(thr), _p(H), None, 0, _p(sc)))
            return H, sc[0], sc[1], sc[2]
        Pb = np.asfortranarray(Pb, dtype=np.float64)
        Hb = np.zeros_like(P, order="F")  # left at zero when beta is false
        check_host(fn(self._h, self._basis._h, int(P.shape[0]), int(x_func), int(c_func), 2, _p(P), _p(Pb), ctypes.c_double(thr), _p(H), _p(Hb),
                      int(bool(beta)), _p(sc)))
        return H, Hb, sc[0], sc[1], sc[2]


    def last_call(self):
        """accounting of the last eval_Fxc: device milliseconds of its kernels (CUDA events), packed bytes up / down, kernel launches"""
        out = np.zeros(3)
        ctx = self._basis.context()
        lib().hfb200_grid_last_call.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        lib().hfb200_grid_launches.argtypes = [ctypes.c_void_p]
        check(lib().hfb200_grid_last_call(ctx, _p(out)))
        return {"device_ms": float(out[0]), "h2d_bytes": int(out[1]), "d2h_bytes": int(out[2]), "launches": int(lib().hfb200_grid_launches(ctx))}


class AtomicDFTGrid(_ProductGrid):
    _prefix = "hfh_atomic"
