"""GPU test of hfb200_sadatom_fock_batch_static, the allocation-free variant of hfb200_sadatom_fock_batch
(scratch in a static device array, include/helfem_b200.h).  Collected last on purpose (file name):
a failure here cannot hide another test.  Self-contained: it builds its own basis and
passes both alone in a fresh process and after the rest of the suite.

Two calls in a row through the C ABI, with different batch sizes and densities (the second call finds the first
call's contents in the shared array), each compared with the existing entry point on the same inputs.  Both run
the same kernels in the same order, so the results must agree to the 1e-12 the batched-API test uses (they are in
fact bit-identical)."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-12


def _call(L, name, ctx, nb, xf, cf, nspin, nsl, dP, dJ, dH, dS, stream):
    return getattr(L, name)(ctx, int(nb), int(xf), int(cf), int(nspin), int(nsl), ctypes.c_double(1e-12),
                            ctypes.c_void_p(dP.data_ptr()), ctypes.c_void_p(dJ.data_ptr() if dJ is not None else 0),
                            ctypes.c_void_p(dH.data_ptr() if dH is not None else 0),
                            ctypes.c_void_p(dS.data_ptr() if dS is not None else 0), ctypes.c_void_p(stream))


def test_static_scratch_entry_matches_fock_batch(native):
    torch = pytest.importorskip("torch")
    import helfem_b200 as hb
    from oracle import fem_oracle as fo
    from util import sadatom_density

    Z, primbas, nnodes, nelem, lmax = 10, 4, 15, 5, 2
    b = hb.SadatomBasis(Z=Z, nelem=nelem, nnodes=nnodes, primbas=primbas, lmax=lmax)
    b.compute_tei(1, False)
    rs = fo.RadialSetup(primbas, nnodes, fo.get_grid(40.0, nelem), taylor_order=-1)
    S, T, Tl, V = rs.overlap(), rs.kinetic(), rs.kinetic_l(), rs.nuclear(Z)
    dens = {s: sadatom_density(S, T, Tl, V, Z, lmax, spin=s) / (4 * np.pi) for s in ("a", "b")}
    nsl, blk = lmax + 1, b.nel * b.np_max**2
    ctx, L = b.context(), native
    stream = torch.cuda.current_stream().cuda_stream

    # (builds, spins, functional pair): restricted PBE with 20 builds (batched Coulomb kernel), then unrestricted
    # r2SCAN with 3 builds (split-K Coulomb kernel) on top of the first call's leftovers
    for nb, nspin, (xf, cf) in ((20, 1, (101, 130)), (3, 2, (497, 498))):
        hostP = np.zeros((nb, nspin, nsl, blk))
        for ib in range(nb):
            for s in range(nspin):
                P = (0.6 + 0.05 * ib) * dens["ab"[s]] if nspin == 2 else (0.6 + 0.05 * ib) * (dens["a"] + dens["b"])
                for l in range(nsl):
                    hostP[ib, s, l] = b.pack_band(P[:, :, l])
        dP = torch.from_numpy(hostP).cuda()
        out = {}
        for name in ("hfb200_sadatom_fock_batch", "hfb200_sadatom_fock_batch_static"):
            dJ = torch.full((nb, blk), np.nan, dtype=torch.float64, device="cuda")
            dH = torch.full((nb, nspin, nsl, blk), np.nan, dtype=torch.float64, device="cuda")
            dS = torch.full((nb, 4), np.nan, dtype=torch.float64, device="cuda")
            torch.cuda.synchronize()
            hb.check(_call(L, name, ctx, nb, xf, cf, nspin, nsl, dP, dJ, dH, dS, stream))
            torch.cuda.synchronize()
            out[name] = (dJ.cpu().numpy(), dH.cpu().numpy(), dS.cpu().numpy())
        ref, new = out["hfb200_sadatom_fock_batch"], out["hfb200_sadatom_fock_batch_static"]
        assert np.all(np.isfinite(ref[0])) and np.max(np.abs(ref[0])) > 1e-3      # the comparison is not vacuous
        assert np.all(np.isfinite(ref[1])) and np.max(np.abs(ref[1])) > 1e-3
        for r, n in zip(ref, new):
            assert np.all(np.isfinite(n))
            assert np.max(np.abs(r - n)) <= TOL * max(1.0, float(np.max(np.abs(r))))
        # Coulomb only (stand-alone pack kernel) and XC only through the static entry
        dJ = torch.full((nb, blk), np.nan, dtype=torch.float64, device="cuda")
        hb.check(_call(L, "hfb200_sadatom_fock_batch_static", ctx, nb, 0, 0, nspin, nsl, dP, dJ, None, None, stream))
        dH = torch.full((nb, nspin, nsl, blk), np.nan, dtype=torch.float64, device="cuda")
        dS = torch.full((nb, 4), np.nan, dtype=torch.float64, device="cuda")
        hb.check(_call(L, "hfb200_sadatom_fock_batch_static", ctx, nb, xf, cf, nspin, nsl, dP, None, dH, dS, stream))
        torch.cuda.synchronize()
        assert np.max(np.abs(dJ.cpu().numpy() - ref[0])) <= TOL * max(1.0, float(np.max(np.abs(ref[0]))))
        assert np.max(np.abs(dH.cpu().numpy() - ref[1])) <= TOL * max(1.0, float(np.max(np.abs(ref[1]))))
        assert np.max(np.abs(dS.cpu().numpy() - ref[2])) <= TOL * max(1.0, float(np.max(np.abs(ref[2]))))

    # a batch beyond the static array is refused with a clear error before anything is launched
    rc = _call(L, "hfb200_sadatom_fock_batch_static", ctx, 200000, 101, 130, 1, nsl, dP, dJ, dH, dS, stream)
    assert rc == hb.EINVAL
    msg = L.hfb200_last_error().decode()
    assert "HFB200_STATIC_SCRATCH_BYTES" in msg and "200000 builds" in msg
