"""ctypes view of hfb::diatomic::TwoDBasis (helfem_b200/host/diatomic.h), the host-side mirror of
helfem::diatomic::basis::TwoDBasis (/root/reference/src/diatomic/basis.h).  Harness only."""
import ctypes

import numpy as np

from . import lib, check, check_host, _p, _ip


class DiatomicBasis:
    """Arguments follow the reference's `diatomic` flags (/root/reference/src/diatomic/main.cpp:89-133):
    lmax is an int (same for every |m| <= mmax) or a list of per-|m| maxima."""

    def __init__(self, Z1, Z2, Rbond, nelem, nnodes=15, primbas=4, nquad=0, rmax=40.0, igrid=4, zexp=2.0, lmax=2, mmax=0,
                 device=0):
        L = lib()
        L.hfh_diatomic_free.argtypes = [ctypes.c_void_p]
        L.hfh_diatomic_free.restype = None
        lm = np.ascontiguousarray([lmax] * (mmax + 1) if np.isscalar(lmax) else lmax, dtype=np.int32)
        self._h = ctypes.c_void_p()
        check_host(L.hfh_diatomic_new(ctypes.byref(self._h), int(Z1), int(Z2), ctypes.c_double(Rbond), int(primbas), int(nnodes),
                                      int(nquad), int(nelem), ctypes.c_double(rmax), int(igrid), ctypes.c_double(zexp),
                                      int(len(lm)), _ip(lm), int(device)))
        v = [ctypes.c_int() for _ in range(6)]
        check_host(L.hfh_diatomic_dims(self._h, *[ctypes.byref(x) for x in v]))
        self.nbf, self.ndummy, self.nrad, self.nang, self.nel, self.nlm = [x.value for x in v]
        self.Rhalf = 0.5 * Rbond
        self.lval = np.zeros(self.nang, dtype=np.int32)
        self.mval = np.zeros(self.nang, dtype=np.int32)
        self.first = np.zeros(self.nel, dtype=np.int32)
        self.last = np.zeros(self.nel, dtype=np.int32)
        self.Lmap = np.zeros(self.nlm, dtype=np.int32)
        self.Mmap = np.zeros(self.nlm, dtype=np.int32)
        check_host(L.hfh_diatomic_tables(self._h, _ip(self.lval), _ip(self.mval), _ip(self.first), _ip(self.last),
                                         _ip(self.Lmap), _ip(self.Mmap)))
        self.have_ktei = False

    def __del__(self):
        try:
            if self._h:
                lib().hfh_diatomic_free(self._h)
                self._h = None
        except Exception:
            pass

    def pure_indices(self):
        idx = []
        for i, m in enumerate(self.mval):
            idx.extend(range(i * self.nrad + (1 if m != 0 else 0), (i + 1) * self.nrad))
        return np.asarray(idx)

    def compute_tei(self, exchange=True):
        check_host(lib().hfh_diatomic_compute_tei(self._h, int(bool(exchange))))
        self.have_ktei = bool(exchange)

    def tei(self, which, ilm, iel):
        """which: 0 P0, 1 P2, 2 Q0, 3 Q2; 4..7 prim_tei 00,02,20,22; 8..11 prim_ktei."""
        npe = int(self.last[iel] - self.first[iel] + 1)
        n = npe if which < 4 else npe * npe
        out = np.zeros((n, n), order="F")
        check_host(lib().hfh_diatomic_get_tei(self._h, int(which), int(ilm), int(iel), _p(out)))
        return out

    def tei_view(self, which, ilm, iel):
        """Same block as tei() WITHOUT a copy (read-only numpy view of the host library's storage; keep the basis alive)."""
        ptr, n = ctypes.POINTER(ctypes.c_double)(), ctypes.c_long()
        check_host(lib().hfh_diatomic_tei_view(self._h, int(which), int(ilm), int(iel), ctypes.byref(ptr), ctypes.byref(n)))
        npe = int(self.last[iel] - self.first[iel] + 1)
        d = npe if which < 4 else npe * npe
        a = np.ctypeslib.as_array(ptr, shape=(n.value,)).reshape((d, d), order="F")
        a.flags.writeable = False
        return a

    def context(self):
        """The hfb200_ctx of this basis (device tables are uploaded on first use)."""
        ctx = ctypes.c_void_p()
        check_host(lib().hfh_diatomic_context(self._h, ctypes.byref(ctx)))
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

    def set_shard(self, rank, nranks, owned_only=False):
        """Multi-GPU: contract only the integral families r % nranks == rank (before the first operator call).
        owned_only: compute_tei builds the stored integrals of the owned families only."""
        check_host(lib().hfh_diatomic_set_shard(self._h, int(rank), int(nranks)))
        check_host(lib().hfh_diatomic_set_owned_only(self._h, int(bool(owned_only))))

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
        check_host(lib().hfh_diatomic_coulomb(self._h, _p(P), _p(J)))

    def exchange_into(self, P, K):
        check_host(lib().hfh_diatomic_exchange(self._h, _p(P), _p(K)))

    def coulomb(self, P):
        P = np.asfortranarray(P, dtype=np.float64)
        J = np.zeros_like(P, order="F")
        check_host(lib().hfh_diatomic_coulomb(self._h, _p(P), _p(J)))
        return J

    def exchange(self, P):
        """returns -K like the reference"""
        P = np.asfortranarray(P, dtype=np.float64)
        K = np.zeros_like(P, order="F")
        check_host(lib().hfh_diatomic_exchange(self._h, _p(P), _p(K)))
        return K

    def radial(self, iel):
        """r (= mu), wrad, bf, df of one element on its quadrature points (radial.get_r / get_wrad / get_bf / get_df,
        /root/reference/src/diatomic/basis.cpp:266-285)."""
        L = lib()
        nq = L.hfh_diatomic_nquad(self._h)
        npe = int(self.last[iel] - self.first[iel] + 1)
        r, w = np.zeros(nq), np.zeros(nq)
        bf, df = np.zeros((nq, npe), order="F"), np.zeros((nq, npe), order="F")
        check_host(L.hfh_diatomic_radial(self._h, int(iel), _p(r), _p(w), _p(bf), _p(df)))
        return r, w, bf, df


from .atomic import _ProductGrid  # noqa: E402


class DiatomicDFTGrid(_ProductGrid):
    """hfb::diatomic::DFTGrid, mirror of helfem::diatomic::dftgrid::DFTGrid (/root/reference/src/diatomic/dftgrid.h:160-187)."""
    _prefix = "hfh_diatomic"
