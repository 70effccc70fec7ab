"""ctypes front-end of tests/hostcheck/hostcheck.cpp (TEST INFRASTRUCTURE: the kernels' shared
scalar header compiled for the host, so its arithmetic can be checked without a GPU)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libhostcheck.so")
_lib = None


_SO_EXACT = os.path.join(_HERE, "libhostcheck_exact.so")
_lib_exact = None


def _build(so, defines=()):
    src = os.path.join(_HERE, "hostcheck.cpp")
    hdr = os.path.join(_HERE, "..", "..", "pythonphot_b200", "csrc", "bphot_math.cuh")
    if (not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr))):
        subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-x", "c++"] + list(defines) +
                              ["-shared", "-o", so, src])
    return C.CDLL(so)


def lib():
    """The header as the product compiles it (default knobs)."""
    global _lib
    if _lib is None:
        _lib = _build(_SO)
    return _lib


def lib_exact():
    """The header with the reference-order arithmetic everywhere (PK_RECIP=0, PW_ONE_ATAN=0): what the bit-for-bit checks use."""
    global _lib_exact
    if _lib_exact is None:
        _lib_exact = _build(_SO_EXACT, ("-DPK_RECIP=0", "-DPW_ONE_ATAN=0"))
    return _lib_exact


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def pixwt(xc, yc, r, x, y, exact=False):
    x = np.ascontiguousarray(x, np.float64); y = np.ascontiguousarray(y, np.float64)
    out = np.empty_like(x)
    (lib_exact() if exact else lib()).hc_pixwt(C.c_double(xc), C.c_double(yc), C.c_double(r), _p(x), _p(y), C.c_int(x.size), _p(out))
    return out


def psfmodel(xx, yy, gauss, psf):
    xx = np.ascontiguousarray(xx, np.float64); yy = np.ascontiguousarray(yy, np.float64)
    gauss = np.ascontiguousarray(gauss, np.float64); psf = np.ascontiguousarray(psf, np.float64)
    out = np.empty((12, xx.size))
    lib().hc_psfmodel(_p(xx), _p(yy), C.c_int(xx.size), _p(gauss), _p(psf), C.c_int(psf.shape[1]), _p(out))
    return out


def inv3(c, exact=False):
    c = np.array(c, dtype=np.float64, order="C")
    st = (lib_exact() if exact else lib()).hc_inv3(_p(c))
    return c, st


def mmm(v, readnoise=0.0, minsky=20, mxiter=50):
    v = np.ascontiguousarray(np.asarray(v).ravel(), np.float64)
    out = np.empty(3)
    it = lib().hc_mmm(_p(v), C.c_int(v.size), C.c_double(readnoise), C.c_int(minsky), C.c_int(mxiter), _p(out))
    return out[0], out[1], out[2], it


def daoerf_separable(xx, yy, gauss):
    xx = np.ascontiguousarray(xx, np.float64); yy = np.ascontiguousarray(yy, np.float64)
    gauss = np.ascontiguousarray(gauss, np.float64)
    return bool(lib().hc_daoerf_separable(_p(xx), _p(yy), C.c_int(xx.size), _p(gauss)))


def mmm_indexed(v, f32=False, readnoise=0.0, minsky=20, mxiter=50):
    """mmm through the index of the sorted vector (what aper_mmm_kernel runs, one thread per star); at most 4096 keys."""
    v = np.ascontiguousarray(np.asarray(v).ravel(), np.float64)
    out = np.empty(3)
    it = lib().hc_mmm_indexed(_p(v), C.c_int(v.size), C.c_int(int(f32)), C.c_double(readnoise), C.c_int(minsky),
                              C.c_int(mxiter), _p(out))
    assert it >= 0
    return out[0], out[1], out[2], it


def indexed_first_ok(v):
    v = np.ascontiguousarray(np.asarray(v).ravel(), np.float64)
    return lib().hc_indexed_first_ok(_p(v), C.c_int(v.size)) == 1
