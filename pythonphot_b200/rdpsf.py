"""Drop-in for PythonPhot/rdpsf.py: read the PSF FITS file written by getpsf and evaluate the full PSF."""
import numpy as np

from . import _fits, dao_value


def rdpsf(psfname, device=None):
    """rdpsf.py:12-67.  Returns ``(psf, hpsf)``: the Gaussian + residual model evaluated on the
    ``(2*psfrad+1)``-pixel grid centred on the star (dao_value on the GPU), and the header as a dict
    with NAXIS1/NAXIS2 set to the size of that grid."""
    resid, hpsf = _fits.read_image(psfname)
    gauss = [hpsf['GAUSS%d' % k] for k in (1, 2, 3, 4, 5)]
    psfrad = hpsf['PSFRAD']
    npsf = int(2 * psfrad + 1)
    d = np.arange(npsf, dtype='int') - psfrad
    dx, dy = np.meshgrid(d, d)
    psf = np.zeros([npsf, npsf]) + np.asarray(
        dao_value.dao_value(dx.reshape(-1), dy.reshape(-1), gauss, resid, deriv=False, device=device)).reshape(npsf, npsf)
    hpsf['NAXIS1'] = npsf
    hpsf['NAXIS2'] = npsf
    return (psf, hpsf)
