"""CPU restatement of PythonPhot/cntrd.py (derivative-search centroid) -- TEST INFRASTRUCTURE ONLY.

Pinned to the unmodified reference by tests/golden/golden_cntrd.npz (tests/test_cntrd.py, CPU leg); used as the
checker for the CUDA cntrd kernel on larger star lists.  Status codes follow include/bphot.h."""
import numpy as np

OK, EDGE, MOVED_EDGE, X_NOT_DECREASING, X_OUT_OF_RANGE, Y_NOT_DECREASING, Y_OUT_OF_RANGE, NO_FINITE_PIXEL = range(8)


def _marginal_shift(d, w, dd, sumc, x_axis):
    """cntrd.py:175-190 / 196-211 for one axis.  ``d``: summed first differences along the axis."""
    sumd = np.sum(w * d)
    sumxd = np.sum(w * dd * d) if x_axis else np.sum(w * d * dd)  # the two axes multiply in a different order
    sumxsq = np.sum(w * dd ** 2)
    if sumxd >= 0:
        return None, True
    return sumxsq * sumd / (sumc * sumxd), False


def cntrd(img, x, y, fwhm, extendbox=0, keepcenter=False):
    """Returns (xcen, ycen, status) arrays; failed stars carry -1, -1 like the reference (cntrd.py:84-231)."""
    img = np.asarray(img)
    ysize, xsize = img.shape
    nhalf = max(int(0.637 * fwhm), 2)  # cntrd.py:91-93
    nbox = 2 * nhalf + 1
    ex, ey = (extendbox, extendbox) if not hasattr(extendbox, "__len__") else extendbox
    ex, ey = int(ex or 0), int(ey or 0)
    x = np.atleast_1d(np.asarray(x, dtype=np.float64))
    y = np.atleast_1d(np.asarray(y, dtype=np.float64))
    xcen, ycen = x.copy(), y.copy()
    status = np.zeros(x.size, dtype=np.int32)
    ix, iy = np.round(x), np.round(y)
    ir = max(nhalf - 1, 1)
    dd = np.arange(nbox - 1).astype(int) + 0.5 - nhalf
    w = 1. - 0.5 * (np.abs(dd) - 0.5) / (nhalf - 0.5)  # cntrd.py:169-170
    sumc = np.sum(w)

    def fail(i, code):
        xcen[i] = ycen[i] = -1
        status[i] = code

    for i in range(x.size):
        if not keepcenter:  # cntrd.py:112-141: brightest pixel of the (extended) box
            hx, hy = nhalf + ex, nhalf + ey
            if ix[i] < hx or ix[i] + hx > xsize - 1 or iy[i] < hy or iy[i] + hy > ysize - 1:
                fail(i, EDGE)
                continue
            big = img[int(iy[i] - hy):int(iy[i] + hy + 1), int(ix[i] - hx):int(ix[i] + hx + 1)]
            finite = big[big == big]
            if finite.size == 0:
                fail(i, NO_FINITE_PIXEL)
                continue
            py, px = np.nonzero(big == np.max(finite))
            if px.size > 1:
                jx, jy = np.round(np.sum(px) / px.size), np.round(np.sum(py) / px.size)
            else:
                jx, jy = px[0], py[0]
            xmax, ymax = ix[i] - hx + jx, iy[i] - hy + jy
        else:
            xmax, ymax = ix[i], iy[i]
        if xmax < nhalf or xmax + nhalf > xsize - 1 or ymax < nhalf or ymax + nhalf > ysize - 1:  # cntrd.py:146-152
            fail(i, MOVED_EDGE)
            continue
        box = img[int(ymax - nhalf):int(ymax + nhalf + 1), int(xmax - nhalf):int(xmax + nhalf + 1)].astype(np.float64)
        # first differences along x, summed over the 2*ir+1 central rows (cntrd.py:175-177)
        dxs = np.sum((box[:, 1:] - box[:, :-1])[nhalf - ir:nhalf + ir + 1, :], 0)
        sh, bad = _marginal_shift(dxs, w, dd, sumc, True)
        if bad:
            fail(i, X_NOT_DECREASING)
            continue
        if np.abs(sh) > nhalf:
            fail(i, X_OUT_OF_RANGE)
            continue
        xc = xmax - sh
        dys = np.sum((box[1:, :] - box[:-1, :])[:, nhalf - ir:nhalf + ir + 1], 1)
        sh, bad = _marginal_shift(dys, w, dd, sumc, False)
        if bad:
            fail(i, Y_NOT_DECREASING)
            continue
        if np.abs(sh) > nhalf:
            fail(i, Y_OUT_OF_RANGE)
            continue
        xcen[i], ycen[i] = xc, ymax - sh
    return xcen, ycen, status
