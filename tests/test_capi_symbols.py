"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/helfem_b200.h declares, and refuses to compute without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "helfem_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(hfb200_[a-z0-9_]+)\s*\(", hdr)))


def test_header_symbols_exported(native):
    syms = declared_symbols()
    assert len(syms) >= 12
    for s in syms:
        assert hasattr(native, s), "libhelfem_b200.so does not export %s" % s


def test_functional_lookup(native):
    import helfem_b200 as hb
    assert hb.xc_id("gga_x_pbe") == 101 and hb.xc_id("mgga_c_r2scan") == 498 and hb.xc_id("lda_c_pw") == 12
    assert hb.parse_method("gga_x_pbe-gga_c_pbe") == (101, 130)
    assert hb.parse_method("lda_x") == (1, 0)
    with pytest.raises(hb.HelfemError) as e:
        hb.xc_id("hyb_gga_xc_b3lyp")
    assert e.value.code == hb.EFUNC


def test_no_cpu_fallback_without_device(native):
    import helfem_b200 as hb
    if hb.device_count() > 0:
        pytest.skip("a GPU is present")
    b = hb.SadatomBasis(Z=2, nelem=2, nnodes=6, lmax=0)
    b.compute_tei(1, False)
    P = np.zeros((b.nbf, b.nbf), order="F")
    with pytest.raises(RuntimeError) as e:
        b.coulomb(P)
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)
    n = 4
    rho = np.ones(n)
    out = [np.zeros(n) for _ in range(4)]
    dp = ctypes.POINTER(ctypes.c_double)
    rc = native.hfb200_xc_eval(0, 1, 12, 1, n, ctypes.c_double(1e-12), rho.ctypes.data_as(dp), None, None,
                               *[o.ctypes.data_as(dp) for o in out])
    assert rc == hb.ENODEVICE


def test_missing_tei_is_logic_error(native):
    import helfem_b200 as hb
    b = hb.SadatomBasis(Z=2, nelem=2, nnodes=6, lmax=0)
    with pytest.raises(ValueError) as e:  # std::logic_error, like sadatom/basis.cpp:297-298
        b.coulomb(np.zeros((b.nbf, b.nbf), order="F"))
    assert "Primitive teis have not been computed" in str(e.value)
