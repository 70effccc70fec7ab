"""pythonphot_b200 -- B200 (sm_100a) batched stellar photometry behind PythonPhot's API.

Module and callable names mirror djones1040/PythonPhot (PythonPhot/__init__.py:14-33):

    from pythonphot_b200 import aper, pkfit, pkfit_noise, pkfit_norecent_noise, mmm, pixwt, dao_value, ...
    mag, magerr, flux, fluxerr, sky, skyerr, badflag, outstr = aper.aper(image, xc, yc, ...)
    pk = pkfit_norecent_noise.pkfit_class(image, gauss, psf, ronois, phpadu, noise, mask)
    errmag, chi, sharp, niter, scale = pk.pkfit_norecent_noise(scale, x, y, sky, radius)

plus the unified ``pkfit_class`` carrying all three fit methods.  Every numerical routine runs as
hand-written CUDA in ``csrc/`` behind the C ABI of ``include/bphot.h``; there is no CPU fallback.
"""
from . import _lib  # noqa: F401
from . import aper, pkfit, pkfit_noise, pkfit_norecent_noise, pkfit_norecenter  # noqa: F401
from . import mmm, pixwt, dao_value, daoerf, rinter  # noqa: F401
from . import getpsf, rdpsf, find, cntrd, shard  # noqa: F401
from ._pkfit_core import PkfitEngine as _PkfitEngine


class pkfit_class(_PkfitEngine):
    """``pkfit_class(image, gauss, psf, ronois, phpadu, noiseim=None, maskim=None)`` with the three fit
    methods the reference spreads over three same-named classes (pkfit.py:119, pkfit_noise.py:118,
    pkfit_norecent_noise.py:119)."""

    pkfit = pkfit.pkfit_class.pkfit
    pkfit_noise = pkfit_noise.pkfit_class.pkfit_noise
    pkfit_norecent_noise = pkfit_norecent_noise.pkfit_class.pkfit_norecent_noise
    pkfit_norecenter = pkfit_norecenter.pkfit_class.pkfit_norecenter


__all__ = ["aper", "pkfit", "pkfit_noise", "pkfit_norecent_noise", "pkfit_norecenter", "mmm", "pixwt", "dao_value", "daoerf",
           "rinter", "getpsf", "rdpsf", "find", "cntrd", "shard", "pkfit_class"]
