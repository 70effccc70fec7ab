"""Drop-in for PythonPhot/pkfit_noise.py (class ``pkfit_class``, method ``pkfit_noise``; pkfit_noise.py:118-462)."""
from . import _lib
from ._pkfit_core import PkfitEngine


class pkfit_class(PkfitEngine):
    """``pkfit_class(image, gauss, psf, ronois, phpadu, noise_image, mask_image)`` (pkfit_noise.py:120-129)."""

    def __init__(self, image, gauss, psf, ronois, phpadu, noise_image, mask_image, device=None):
        PkfitEngine.__init__(self, image, gauss, psf, ronois, phpadu, noise_image, mask_image, device=device)

    def pkfit_noise(self, scale, x, y, sky, radius, maxiter=25, debug=False, debug2=False, xyout=False):
        """Recentring PSF fit weighted by the noise image, masked pixels excluded (pkfit_noise.py:131-462).

        Scalars (reference usage) or equal-length arrays (batched).  Returns
        ``(errmag, chi, sharp, niter, scale)`` plus ``(x, y)`` when ``xyout``."""
        return self.fit(_lib.PKFIT_NOISE, scale, x, y, sky, radius, maxiter=maxiter, xyout=xyout)
