"""CPU check of the device XC source: helfem_b200/csrc/xc_point.cuh compiled for the host by a TEST-ONLY
shim (tests/xc_host_shim.cpp; never part of the product library) against the oracle's independent
complex-step restatement, functional by functional.  The GPU twin is
tests/test_sadatom_gpu.py::test_xc_pointwise_device_vs_oracle."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

import oracle
from util import xc_pointwise_check

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def shim():
    so = os.path.join(HERE, "_xc_host.so")
    src = os.path.join(HERE, "xc_host_shim.cpp")
    deps = [src] + [os.path.join(HERE, "..", "helfem_b200", "csrc", f) for f in ("xc_point.cuh", "xc_functionals.cuh", "dual.cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-x", "c++", "-o", so, src])
    return ctypes.CDLL(so)


def test_device_xc_source_matches_oracle_on_host(shim):
    dp = ctypes.POINTER(ctypes.c_double)

    def evaluate(f, nspin, n, r_, s_, t_):
        exc, vr, vs, vt = np.zeros(n), np.zeros_like(r_), np.zeros_like(s_), np.zeros_like(t_)
        is_x = f in (1, 101, 497)
        rc = shim.xc_host_eval(f if is_x else 0, 0 if is_x else f, nspin, n, ctypes.c_double(1e-12), r_.ctypes.data_as(dp),
                               s_.ctypes.data_as(dp), t_.ctypes.data_as(dp), exc.ctypes.data_as(dp), vr.ctypes.data_as(dp),
                               vs.ctypes.data_as(dp), vt.ctypes.data_as(dp))
        assert rc == 0
        return exc, vr, vs, vt

    xc_pointwise_check(evaluate, oracle.xc)  # asserts the gate per functional / spin / output


def test_threshold_returns_exact_zeros(shim):
    dp = ctypes.POINTER(ctypes.c_double)
    n = 8
    rho = np.full(n, 1e-13)
    sg, ta = np.full(n, 1e-30), np.full(n, 1e-20)
    for x, c in ((1, 12), (101, 130), (497, 498)):
        exc, vr, vs, vt = np.ones(n), np.ones(n), np.ones(n), np.ones(n)
        shim.xc_host_eval(x, c, 1, n, ctypes.c_double(1e-12), rho.ctypes.data_as(dp), sg.ctypes.data_as(dp), ta.ctypes.data_as(dp),
                          exc.ctypes.data_as(dp), vr.ctypes.data_as(dp), vs.ctypes.data_as(dp), vt.ctypes.data_as(dp))
        assert not exc.any() and not vr.any()
