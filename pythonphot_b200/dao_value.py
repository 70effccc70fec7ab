"""Drop-in for PythonPhot/dao_value.py: value of a DAOPHOT PSF (Gaussian + residual table) at points."""
import numpy as np

from . import _psfmodel


def dao_value(xx, yy, gauss, psf, deriv=False, device=None):
    """dao_value.py:9-121.  ``xx, yy`` are offsets from the PSF centre (1-D arrays or scalars).

    Returns ``value`` or ``(value, dvdx, dvdy)``.  If ANY point falls off the look-up table the
    reference prints a message and returns zeros for ALL points (dao_value.py:70-84); so does this."""
    xa = np.atleast_1d(np.asarray(xx, dtype=np.float64)).reshape(-1)
    ya = np.atleast_1d(np.asarray(yy, dtype=np.float64)).reshape(-1)
    npsf = np.shape(psf)[1]
    half = float(npsf - 1) / 2.
    x, y = 2. * xa + half, 2. * ya + half
    if xa.size == 0 or (x.min() < 1.) or (x.max() > npsf - 2.) or (y.min() < 1.) or (y.max() > npsf - 2.):
        print('X,Y positions too close to edge of frame' if xa.size else 'No coordinates provided')
        z = xa * 0
        return (z, z.copy(), z.copy()) if deriv else z
    o = _psfmodel.evaluate(xa, ya, gauss, psf, device=device)
    return (o[0], o[1], o[2]) if deriv else o[0]
