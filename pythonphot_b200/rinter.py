"""Drop-in for PythonPhot/rinter.py: bicubic (IDL CUBIC=-0.5) interpolation of a table with derivatives."""
import numpy as np

from . import _lib, _psfmodel

_G = np.array([1.0, 0.0, 0.0, 1.0, 1.0])


def rinter(p, x, y, deriv=True, device=None):
    """rinter.py:8-214 for 1-D point lists on a SQUARE table (the PSF look-up table use case).

    ``x, y`` are table coordinates (x = column).  Returns ``z`` or ``(z, dfdx, dfdy)``.  Points must lie in
    [1, n-2] (where dao_value.py:70-84 guarantees them); outside that range the reference reads wrapped or
    padded flat indices (rinter.py:117-157), which is not reproduced here."""
    p = np.asarray(p, dtype=np.float64)
    if p.ndim != 2 or p.shape[0] != p.shape[1]:
        raise NotImplementedError("rinter: only square 2-D tables are supported on the GPU path")
    n = p.shape[1]
    xa = np.atleast_1d(np.asarray(x, dtype=np.float64)).reshape(-1)
    ya = np.atleast_1d(np.asarray(y, dtype=np.float64)).reshape(-1)
    if xa.size and ((xa.min() < 1.) or (xa.max() > n - 2.) or (ya.min() < 1.) or (ya.max() > n - 2.)):
        raise _lib.BphotError("rinter: points outside [1, n-2] are not supported")
    o = _psfmodel.evaluate(xa, ya, _G, p, device=device, table_coords=True)
    return (o[9], o[10], o[11]) if deriv else o[9]
