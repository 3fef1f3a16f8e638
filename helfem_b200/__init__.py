"""helfem_b200 -- B200-native Fock build (J + K + Vxc, FP64) behind HelFEM's operator interfaces.

The product is the native library ``libhelfem_b200.so`` (C ABI in ``include/helfem_b200.h``: CUDA
kernels for sm_100a plus the C++ host-side mirror of the reference's ``TwoDBasis`` / ``DFTGrid``
classes).  This Python package is only the harness-side binding used by the tests and by
``bench.py``: ctypes prototypes and thin numpy wrappers.  It never computes anything itself and
it fails loudly when the native library is missing -- there is no Python or CPU fallback.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# HELFEM_B200_LIB: a tuning variant built by tools/build_variant.py (A/B runs); default = the in-tree build
LIB_PATH = os.environ.get("HELFEM_B200_LIB") or os.path.join(_HERE, "libhelfem_b200.so")

_lib = None

c_double_p = ctypes.POINTER(ctypes.c_double)
c_int_p = ctypes.POINTER(ctypes.c_int)


class NativeLibraryMissing(RuntimeError):
    pass


class HelfemError(RuntimeError):
    """Non-zero status from the C ABI (message = hfb200_last_error())."""

    def __init__(self, code, msg):
        super().__init__("helfem_b200 status %d: %s" % (code, msg))
        self.code = code


# status codes of include/helfem_b200.h
OK, EINVAL, ENOTEI, EFUNC, ENODEVICE, ECUDA = 0, -1, -2, -3, -4, -5

XC_IDS = {"none": 0, "lda_x": 1, "lda_c_pw": 12, "lda_c_pw_mod": 13, "gga_x_pbe": 101, "gga_c_pbe": 130,
          "mgga_x_r2scan": 497, "mgga_c_r2scan": 498}


def lib():
    """Load libhelfem_b200.so (built in-tree by ``python -m helfem_b200.build``)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NativeLibraryMissing(
                "%s not found: build it with `python -m helfem_b200.build` (or __graft_entry__.build()); "
                "helfem_b200 has no fallback implementation" % LIB_PATH)
        _lib = ctypes.CDLL(LIB_PATH)
        _declare(_lib)
    return _lib


def _declare(L):
    L.hfb200_last_error.restype = ctypes.c_char_p
    L.hfb200_version.restype = ctypes.c_char_p
    L.hfb200_xc_id.argtypes = [ctypes.c_char_p]
    L.hfh_last_error.restype = ctypes.c_char_p
    L.hfh_gaunt.restype = ctypes.c_double
    L.hfh_gaunt.argtypes = [ctypes.c_int] * 6
    L.hfh_modified_gaunt.restype = ctypes.c_double
    L.hfh_modified_gaunt.argtypes = [ctypes.c_int] * 6
    L.hfb200_destroy.argtypes = [ctypes.c_void_p]
    L.hfb200_destroy.restype = None
    L.hfh_sadatom_free.argtypes = [ctypes.c_void_p]
    L.hfh_sadatom_free.restype = None


def _p(a):
    return a.ctypes.data_as(c_double_p)


def _ip(a):
    return a.ctypes.data_as(c_int_p)


def check(rc):
    if rc != 0:
        raise HelfemError(rc, lib().hfb200_last_error().decode())


def check_host(rc):
    """hostapi.cpp status: 1 = std::logic_error, 2 = std::runtime_error."""
    if rc != 0:
        msg = lib().hfh_last_error().decode()
        if rc == 1:
            raise ValueError("std::logic_error: " + msg)
        raise RuntimeError("std::runtime_error: " + msg)


def xc_id(name):
    r = lib().hfb200_xc_id(name.encode())
    if r < 0:
        raise HelfemError(EFUNC, lib().hfb200_last_error().decode())
    return r


def parse_method(method):
    """'gga_x_pbe-gga_c_pbe' -> (x_id, c_id), exchange first as in general/dftfuncs.cpp:94-117."""
    parts = method.split("-")
    if len(parts) == 1:
        name = parts[0]
        if "_x_" in name or name.endswith("_x"):
            return xc_id(name), 0
        return 0, xc_id(name)
    return xc_id(parts[0]), xc_id(parts[1])


def comm_unique_id():
    """NCCL unique id (bytes) for hfb200_comm_init; call on rank 0 and hand it to the other ranks."""
    buf = ctypes.create_string_buffer(128)
    check(lib().hfb200_comm_unique_id(buf, 128))
    return buf.raw


def device_count():
    return lib().hfb200_device_count()


from .sadatom import SadatomBasis  # noqa: E402,F401
from .atomic import AtomicBasis, AtomicDFTGrid  # noqa: E402,F401
from .diatomic import DiatomicBasis, DiatomicDFTGrid  # noqa: E402,F401
