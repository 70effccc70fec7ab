"""Drop-in for PythonPhot/pkfit_norecenter.py (class ``pkfit_class``, method ``pkfit_norecenter``;
pkfit_norecenter.py:119-657), the fit ``photfunctions.py:340-372`` calls on fake stars."""
from . import _lib
from ._pkfit_core import PkfitEngine


class pkfit_class(PkfitEngine):
    """``pkfit_class(image, gauss, psf, ronois, phpadu, weightim=None)`` (pkfit_norecenter.py:120-126; the weight image is
    only read by ``pkfit_fast_norecenter``)."""

    def __init__(self, image, gauss, psf, ronois, phpadu, weightim=None, device=None):
        PkfitEngine.__init__(self, image, gauss, psf, ronois, phpadu, None, None, device=device)
        self.w = weightim

    def pkfit_norecenter(self, scale, x, y, sky, radius, debug=False, maxiter=25):
        """Scale-only PSF fit at a frozen position with pkfit.py's analytic noise model (pkfit_norecenter.py:346-657).

        The reference method raises under Python 3 (``dy[good / ixx]``, pkfit_norecenter.py:435); what runs here is the
        method under the Python-2 division it was written for, pinned to goldens generated that way
        (tests/golden/gen_golden_norecenter.py).  It is not ``pkfit(recenter=False)``: the all-pixels-rejected return, the
        NaN-matrix branch and the convergence test differ.  Scalars or equal-length arrays; returns
        ``(errmag, chi, sharp, niter, scale)``."""
        return self.fit(_lib.PKFIT_NORECENTER, scale, x, y, sky, radius, maxiter=maxiter, recenter=False)

    def pkfit_fast_norecenter(self, *args, **kw):
        """pkfit_norecenter.py:128-344 (a ``scipy.optimize.leastsq`` wrapper on the host) is outside the accelerated
        path."""
        raise NotImplementedError("pkfit_fast_norecenter is a scipy.optimize.leastsq host routine outside the accelerated path; "
                                  "use pkfit_norecenter")
