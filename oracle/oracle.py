"""ctypes front-end of the CPU oracle (oracle/bphot_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  Never by the product package.
Function names mirror the reference modules they restate.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libbphot_oracle.so")
_lib = None

PKFIT, PKFIT_NOISE, PKFIT_NORECENT, PKFIT_NORECENTER = 0, 1, 2, 3


def build(force=False):
    src = os.path.join(_HERE, "bphot_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "libbphot_oracle.so"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.bo_pixwt.restype = C.c_double
        _lib.bo_pixwt.argtypes = [C.c_double] * 5
        _lib.bo_daoerf.restype = C.c_double
        _lib.bo_rinter.restype = C.c_double
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _img(a):
    a = np.asarray(a)
    if a.dtype == np.float32:
        return np.ascontiguousarray(a), 0
    return np.ascontiguousarray(a, dtype=np.float64), 1


def max_threads():
    return lib().bo_max_threads()


def Pixwt(xc, yc, r, x, y):
    x = np.ascontiguousarray(np.atleast_1d(x), dtype=np.float64)
    y = np.ascontiguousarray(np.atleast_1d(y), dtype=np.float64)
    out = np.empty_like(x)
    lib().bo_pixwt_vec(C.c_double(xc), C.c_double(yc), C.c_double(r), _p(x), _p(y), C.c_int(x.size), _p(out))
    return out


def mmm(sky_vector, readnoise=0.0, minsky=20, mxiter=50, return_info=False):
    v = np.ascontiguousarray(np.asarray(sky_vector).ravel(), dtype=np.float64)
    work = np.empty(max(v.size, 1))
    mod, sig, skw = C.c_double(), C.c_double(), C.c_double()
    info = (C.c_int * 2)()
    lib().bo_mmm.restype = C.c_int
    it = lib().bo_mmm(_p(v), C.c_int(v.size), C.c_double(float(readnoise or 0.0)), C.c_int(minsky), C.c_int(mxiter),
                      C.byref(mod), C.byref(sig), C.byref(skw), info, _p(work))
    if return_info:
        return mod.value, sig.value, skw.value, it, info[0], info[1]
    return mod.value, sig.value, skw.value


def aper(image, xc, yc, phpadu=1, apr=5, zeropoint=25, skyrad=(40, 50), badpix=(0, 0),
         setskyval=None, minsky=20, exact=False, readnoise=0, compat="py3", nthreads=1,
         return_extra=False):
    """Same argument meaning as aper.py:26-29; always returns [nstars,naper] arrays."""
    img, dt = _img(image)
    xc = np.ascontiguousarray(np.atleast_1d(xc), dtype=np.float64)
    yc = np.ascontiguousarray(np.atleast_1d(yc), dtype=np.float64)
    apr = np.ascontiguousarray(np.atleast_1d(apr), dtype=np.float64)
    n, na = xc.size, apr.size
    outs = [np.empty((n, na)) for _ in range(4)]
    bad = np.empty((n, na))
    sky, skyerr = np.empty(n), np.empty(n)
    nsky, mit = np.empty(n, np.int32), np.empty(n, np.int32)
    fragile = np.zeros((n, na), np.uint8)
    alt5 = np.full((n, na, 5), np.nan)
    if setskyval is not None:
        ssv = np.atleast_1d(np.asarray(setskyval, dtype=np.float64))
        if ssv.size == 1:
            ssv = np.array([ssv[0], 0.0, 1.0])
        use = 1
    else:
        ssv, use = np.zeros(3), 0
    if minsky is None or (hasattr(minsky, "__len__") and len(minsky) == 0):
        minsky = 20
    lib().bo_aper(_p(img), C.c_int(dt), C.c_int(img.shape[0]), C.c_int(img.shape[1]), _p(xc), _p(yc), C.c_int(n),
                  C.c_double(phpadu), _p(apr), C.c_int(na), C.c_double(zeropoint),
                  C.c_double(skyrad[0]), C.c_double(skyrad[1]), C.c_double(badpix[0]), C.c_double(badpix[1]),
                  C.c_int(use), _p(ssv), C.c_int(int(minsky)), C.c_int(int(bool(exact))), C.c_double(float(readnoise or 0)),
                  C.c_int(1 if compat == "py2" else 0),
                  _p(outs[0]), _p(outs[1]), _p(outs[2]), _p(outs[3]), _p(sky), _p(skyerr), _p(bad), _p(nsky), _p(mit),
                  _p(fragile), _p(alt5), C.c_int(nthreads))
    res = (outs[0], outs[1], outs[2], outs[3], sky, skyerr, bad)
    if return_extra:
        # [9] fragile (bool), [10] the admissible alternative outcome of fragile apertures (oracle/compare.py)
        alt = {"flux": alt5[..., 0], "fluxerr": alt5[..., 1], "mag": alt5[..., 2], "magerr": alt5[..., 3], "badflag": alt5[..., 4],
               "kind": fragile.copy(), "phpadu": float(phpadu), "zeropoint": float(zeropoint), "apr": apr.copy(), "nsky": nsky.copy(),
               "setsky": (ssv.copy() if use else None)}
        return res + (nsky, mit, fragile.astype(bool), alt)
    return res


def daoerf(x, y, a):
    x = np.atleast_1d(np.asarray(x, float)).ravel()
    y = np.atleast_1d(np.asarray(y, float)).ravel()
    a = np.ascontiguousarray(a, dtype=np.float64)
    f = np.empty(x.size)
    pder = np.empty((5, x.size))
    tmp = (C.c_double * 5)()
    for i in range(x.size):
        f[i] = lib().bo_daoerf(C.c_double(x[i]), C.c_double(y[i]), _p(a), tmp)
        pder[:, i] = tmp[:]
    return f, pder


def rinter(p, x, y):
    p = np.ascontiguousarray(p, dtype=np.float64)
    x = np.atleast_1d(np.asarray(x, float)).ravel()
    y = np.atleast_1d(np.asarray(y, float)).ravel()
    z, dx, dy = np.empty(x.size), np.empty(x.size), np.empty(x.size)
    a, b = C.c_double(), C.c_double()
    for i in range(x.size):
        z[i] = lib().bo_rinter(_p(p), C.c_int(p.shape[0]), C.c_int(p.shape[1]), C.c_double(x[i]), C.c_double(y[i]),
                               C.byref(a), C.byref(b))
        dx[i], dy[i] = a.value, b.value
    return z, dx, dy


def dao_value(xx, yy, gauss, psf):
    xx = np.ascontiguousarray(np.atleast_1d(xx), dtype=np.float64)
    yy = np.ascontiguousarray(np.atleast_1d(yy), dtype=np.float64)
    gauss = np.ascontiguousarray(gauss, dtype=np.float64)
    psf = np.ascontiguousarray(psf, dtype=np.float64)
    v, dx, dy = np.empty(xx.size), np.empty(xx.size), np.empty(xx.size)
    lib().bo_dao_value(_p(xx), _p(yy), C.c_int(xx.size), _p(gauss), _p(psf), C.c_int(psf.shape[1]), _p(v), _p(dx), _p(dy))
    return v, dx, dy


def inv3(c):
    c = np.array(c, dtype=np.float64, order="C")
    st = C.c_int()
    lib().bo_inv3(_p(c), C.byref(st))
    return c, st.value


def pkfit(variant, image, gauss, psf, ronois, phpadu, noise, mask, scale, x, y, sky, radius,
          maxiter=25, recenter=True, nthreads=1):
    """Batched restatement of pkfit_class(...).pkfit / pkfit_noise / pkfit_norecent_noise.

    Returns dict(errmag, chi, sharp, niter, scale, x, y, chi2), arrays of len(x)."""
    img, dt = _img(image)
    if variant in (PKFIT, PKFIT_NORECENTER):
        nz, ndt, mk = img, dt, np.zeros(1, np.uint8)
    else:
        nz, ndt = _img(noise)
        mk = np.ascontiguousarray(np.asarray(mask) != 0, dtype=np.uint8)
    gauss = np.ascontiguousarray(gauss, dtype=np.float64)
    psf = np.ascontiguousarray(psf, dtype=np.float64)
    x = np.ascontiguousarray(np.atleast_1d(x), dtype=np.float64)
    n = x.size
    y = np.ascontiguousarray(np.broadcast_to(np.asarray(y, np.float64), (n,)))
    scale = np.ascontiguousarray(np.broadcast_to(np.asarray(scale, np.float64), (n,)))
    sky = np.ascontiguousarray(np.broadcast_to(np.asarray(sky, np.float64), (n,)))
    o = {k: np.empty(n) for k in ("errmag", "chi", "sharp", "scale", "x", "y", "chi2")}
    o["niter"] = np.empty(n, np.int32)
    lib().bo_pkfit(C.c_int(variant), _p(img), C.c_int(dt), _p(nz), C.c_int(ndt), _p(mk),
                   C.c_int(img.shape[0]), C.c_int(img.shape[1]), _p(gauss), _p(psf), C.c_int(psf.shape[1]),
                   C.c_double(ronois), C.c_double(phpadu), _p(scale), _p(x), _p(y), _p(sky),
                   C.c_double(radius), C.c_int(maxiter), C.c_int(int(bool(recenter))), C.c_int(n),
                   _p(o["errmag"]), _p(o["chi"]), _p(o["sharp"]), _p(o["niter"]), _p(o["scale"]),
                   _p(o["x"]), _p(o["y"]), _p(o["chi2"]), C.c_int(nthreads))
    return o
