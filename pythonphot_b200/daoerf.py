"""Drop-in for PythonPhot/daoerf.py: pixel-integrated bivariate Gaussian and its 5 partial derivatives."""
import numpy as np

from . import _psfmodel

_DUMMY = np.zeros((4, 4))


def daoerf(x, y, a, device=None):
    """daoerf.py:9-74.  Returns ``(f, pder)`` with ``pder`` of shape [5, npts]."""
    xa = np.atleast_1d(np.asarray(x, dtype=np.float64)).reshape(-1)
    ya = np.atleast_1d(np.asarray(y, dtype=np.float64)).reshape(-1)
    o = _psfmodel.evaluate(xa, ya, a, _DUMMY, device=device)
    return o[3], o[4:9].copy()
