"""Drop-in for PythonPhot/pkfit_norecent_noise.py (class ``pkfit_class``, method
``pkfit_norecent_noise``; pkfit_norecent_noise.py:119-531)."""
from . import _lib
from ._pkfit_core import PkfitEngine


class pkfit_class(PkfitEngine):
    """``pkfit_class(image, gauss, psf, ronois, phpadu, noise_image, mask_image)``
    (pkfit_norecent_noise.py:121-130)."""

    def __init__(self, image, gauss, psf, ronois, phpadu, noise_image, mask_image, device=None):
        PkfitEngine.__init__(self, image, gauss, psf, ronois, phpadu, noise_image, mask_image, device=device)

    def pkfit_norecent_noise(self, scale, x, y, sky, radius, maxiter=25, debug=False, returnchi2=False,
                             verbose=False):
        """Forced-position PSF fit (scale only; x, y frozen) weighted by the noise image
        (pkfit_norecent_noise.py:132-531).

        Scalars (reference usage) or equal-length arrays (batched).  Returns
        ``(errmag, chi, sharp, niter, scale)`` plus ``chi2`` when ``returnchi2``.  (The reference drops
        chi2 from the tuple on its NaN-position and infinite-matrix early returns,
        pkfit_norecent_noise.py:156-164, 462-468; here the tuple length never changes, chi2 is NaN.)"""
        return self.fit(_lib.PKFIT_NORECENT, scale, x, y, sky, radius, maxiter=maxiter, returnchi2=returnchi2)
