"""Drop-in for PythonPhot/rinter.py: bicubic (IDL CUBIC=-0.5) interpolation of a table with derivatives."""
import numpy as np

from . import _lib, _psfmodel

_G = np.array([1.0, 0.0, 0.0, 1.0, 1.0])


def rinter(p, x, y, deriv=True, device=None):
    """rinter.py:8-214 on a SQUARE table.  ``x, y`` are table coordinates (x = column), any shape (flattened).

    Returns ``z`` or ``(z, dfdx, dfdy)``.  The reference indexes the FLATTENED table (rinter.py:101-104): rows wrap
    into each other at the table edges, flat indices past the end read the last element (its padded copy,
    rinter.py:117-135) and negative ones wrap like any numpy index.  All of that is reproduced; only a negative
    index that numpy itself would reject raises."""
    p = np.asarray(p, dtype=np.float64)
    if p.ndim != 2 or p.shape[0] != p.shape[1]:
        raise NotImplementedError("rinter: only square 2-D tables are supported on the GPU path")
    n = p.shape[1]
    xa = np.atleast_1d(np.asarray(x, dtype=np.float64)).reshape(-1)
    ya = np.atleast_1d(np.asarray(y, dtype=np.float64)).reshape(-1)
    mode = 1
    if xa.size:
        km = n * (ya.astype(int) - 1) + xa.astype(int)  # rinter.py:98-104 (truncation)
        kmin, kmax = int(km.min()) - 1, int(km.max()) + 3 * n
        if kmin < 0:
            wrap = kmax + 3 if kmax > p.size - 3 else p.size  # the array the reference indexes: padded copy or the table
            if kmin + wrap < 0:
                raise IndexError("rinter: flat index %d is out of bounds for the table" % kmin)
            mode = int(wrap)
    o = _psfmodel.evaluate(xa, ya, _G, p, device=device, table_coords=mode)
    return (o[9], o[10], o[11]) if deriv else o[9]
