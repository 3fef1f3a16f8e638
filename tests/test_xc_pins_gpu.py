"""GPU twin of tests/test_xc_pins.py (run with -m gpu): the DEVICE evaluation of the six functionals (hfb200_xc_eval through the
C ABI) against the committed golden vectors of the independent 110-digit mpmath restatement (tests/golden/xc_mpmath.npz),
1024 seeded points, polarised and unpolarised, every libxc-style output, gate 1e-11 relative."""
import ctypes

import numpy as np
import pytest

from test_xc_pins import compare_with_goldens

pytestmark = pytest.mark.gpu


def test_device_xc_matches_mpmath_goldens(native):
    dp = ctypes.POINTER(ctypes.c_double)

    def evaluate(f, nspin, rho, sigma, tau):
        r_, s_, t_ = (np.ascontiguousarray(a, dtype=np.float64).ravel().copy() for a in (rho, sigma, tau))
        n = r_.size // nspin
        exc, vr, vs, vt = np.zeros(n), np.zeros_like(r_), np.zeros_like(s_), np.zeros_like(t_)
        is_x = f in (1, 101, 497)
        rc = native.hfb200_xc_eval(0, f if is_x else 0, 0 if is_x else f, nspin, n, ctypes.c_double(1e-12), r_.ctypes.data_as(dp),
                                   s_.ctypes.data_as(dp), t_.ctypes.data_as(dp), exc.ctypes.data_as(dp), vr.ctypes.data_as(dp),
                                   vs.ctypes.data_as(dp), vt.ctypes.data_as(dp))
        assert rc == 0
        return exc, vr, vs, vt

    worst = compare_with_goldens(evaluate)
    print("device XC vs mpmath goldens: worst relative deviation %.2e" % worst)
    assert worst < 1e-11
