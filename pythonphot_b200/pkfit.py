"""Drop-in for PythonPhot/pkfit.py (class ``pkfit_class``, method ``pkfit``; pkfit.py:119-458)."""
from . import _lib
from ._pkfit_core import PkfitEngine


class pkfit_class(PkfitEngine):
    """``pkfit_class(image, gauss, psf, ronois, phpadu, noiseim=None, maskim=None)`` (pkfit.py:121-130).

    The analytic-noise fit never reads noiseim/maskim.  (The reference raises ValueError at
    pkfit.py:203 when arrays are passed for them; here they are accepted and ignored by ``pkfit``.)"""

    def pkfit(self, scale, x, y, sky, radius, debug=False, xyout=False, maxiter=25, recenter=True):
        """Damped Gauss-Newton PSF fit of (scale, x, y) with the analytic noise model (pkfit.py:132-458).

        ``scale, x, y, sky`` may be scalars (reference usage) or equal-length arrays (batched).
        Returns ``(errmag, chi, sharp, niter, scale)`` plus ``(x, y)`` when ``xyout``."""
        return self.fit(_lib.PKFIT, scale, x, y, sky, radius, maxiter=maxiter, recenter=recenter, xyout=xyout)
