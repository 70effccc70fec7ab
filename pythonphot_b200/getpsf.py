"""Drop-in for PythonPhot/getpsf.py: build the DAOPHOT PSF model (Gaussian + residual table) from stars.

getpsf is sequential by construction -- star k is fitted with the table accumulated from stars
0..k-1 (getpsf.py:308-386) -- and touches a few tens of stars, so it stays host orchestration
 over the same CUDA primitives the photometry path uses:

  * the integrated Gaussian and its derivatives  -> ``daoerf``  (csrc psfmodel_kernel, daoerf.py:49-74)
  * residual interpolation onto the half-pixel grid -> ``rinter`` (rinter.py:98-195)
  * the star-by-star profile fit                   -> ``pkfit.pkfit_class.pkfit`` (pkfit_kernel, batch of 1)

Only the 5x5 normal-equation solve of the first star's Gaussian fit (25 numbers) and the running
sums of the table run in numpy.  The PSF FITS file is written by ``_fits`` (no astropy needed).
"""
import numpy as np
from scipy import linalg

from . import _fits, daoerf, rinter, pkfit


def _box_bounds(xstar, ystar, nhalf):
    # getpsf.py:175-176: NBOX-wide window with the centroid between NBOX/2 and NBOX/2+1
    ix, iy = int(xstar), int(ystar)
    return ix - nhalf + 1, ix + nhalf, iy - nhalf + 1, iy + nhalf


def _fit_core_gaussian(f, x, y, device):
    """Five-parameter integrated-Gaussian fit to the core of the first PSF star (getpsf.py:186-283).

    ``f`` is the sky-subtracted NBOX x NBOX window, (x, y) the centroid inside it.  Returns
    ``(h, dxcen, dycen, sigx, sigy)`` or None when 100 iterations did not converge."""
    ix, iy = int(x + 0.5), int(y + 0.5)
    h = np.max(f)
    sigx = sigy = 2.0
    dxcen = dycen = 0.0
    niter = 0
    while True:
        niter += 1
        if niter > 100:  # getpsf.py:227-233
            return None
        nx = 1 if sigx <= 1 else (3 if sigx > 3 else 2)  # half-width of the fitting window per axis
        ny = 1 if sigy <= 1 else (3 if sigy > 3 else 2)
        cols = np.arange(ix - nx, ix + nx + 1, dtype=np.float64)
        rows = np.arange(iy - ny, iy + ny + 1, dtype=np.float64)
        gx, gy = np.meshgrid(cols, rows)
        model, t = daoerf.daoerf(gx, gy, np.array([h, x + dxcen, y + dycen, sigx, sigy]), device=device)
        resid = (f[iy - ny:iy + ny + 1, ix - nx:ix + nx + 1] - model.reshape(2 * ny + 1, 2 * nx + 1)).reshape(-1)
        v = np.array([np.sum(resid * t[k]) for k in range(5)])
        c = np.array([[np.sum(t[k] * t[l]) for k in range(5)] for l in range(5)])  # c[l, k], getpsf.py:263
        try:
            cinv = linalg.inv(c)
        except Exception:  # singular: the reference repeats the iteration unchanged until the 100 cap
            continue
        z = np.asarray(np.matmul(v.reshape(1, 5), cinv))[0]
        h = h + z[0] / (1.0 + 4.0 * np.abs(z[0] / h))
        dxcen = dxcen + z[1] / (1.0 + 3.0 * np.abs(z[1]))
        dycen = dycen + z[2] / (1.0 + 3.0 * np.abs(z[2]))
        sigx = sigx + z[3] / (1.0 + 4.0 * np.abs(z[3] / sigx))
        sigy = sigy + z[4] / (1.0 + 4.0 * np.abs(z[4] / sigy))
        if np.abs(z[0] / h) + np.abs(z[3] / sigx) + np.abs(z[4] / sigy) < 0.0001:
            return h, dxcen, dycen, sigx, sigy


def _patch_nans(tab):
    """NaN entries of an interpolated residual table -> median of the finite entries in the 5x5 window
    around them, else of the whole table (getpsf.py:357-373).  Patched values feed later windows, in
    row-major order, as in the reference."""
    bad = np.argwhere(tab != tab)
    nr, nc = tab.shape
    for r, c in bad:
        sub = tab[max(r - 2, 0):min(r + 3, nr), max(c - 2, 0):min(c + 3, nc)]
        tab[r, c] = np.median(sub[sub == sub])  # an all-NaN window gives NaN (np.median of nothing), as the reference
    return tab


def getpsf(image, xc, yc,
           apmag, sky, ronois,
           phpadu, idpsf, psfrad,
           fitrad, psfname, zeropoint=0,
           debug=False, verbose=True, device=None):
    """Generate a PSF from observed stars (reference: getpsf.py:15-427; same arguments, same return).

    Returns ``(gauss[5], psf[npsf, npsf], psfmag)``; ``(gauss, -1, -1)`` on failure (fitrad >= psfrad,
    no usable star, no convergence on the last candidate).  ``psfname`` = None or '' skips the FITS
    file.  Where the reference would raise NameError (every candidate first star too close to the
    edge, getpsf.py:177-179 then :334), this returns the failure tuple."""
    image = np.asarray(image)
    nrow, ncol = image.shape
    xc, yc = np.atleast_1d(np.asarray(xc, dtype=np.float64)), np.atleast_1d(np.asarray(yc, dtype=np.float64))
    sky = np.atleast_1d(np.asarray(sky, dtype=np.float64))
    mag = np.asarray(apmag, dtype=np.float64)
    if mag.ndim > 1:  # several apertures: use the first (getpsf.py:61-63)
        mag = mag[:, 0]
    mag = np.atleast_1d(mag)
    psfmag = 1000000.0
    gauss = np.array([1, -1, -1, -1, -1])
    if fitrad >= psfrad:
        print('ERROR - Fitting radius must be smaller than radius defining PSF')
        return (gauss, -1, -1)
    idpsf = list(np.atleast_1d(idpsf))
    numpsf = len(idpsf)
    n = 2 * int(psfrad + 0.5) + 1  # getpsf.py:142-145
    npsf = 2 * n + 7
    nbox = n + 7
    nhalf = nbox / 2.
    gen = np.arange(nbox, dtype=np.float64)
    xgen, ygen = np.meshgrid(gen, gen)
    psfgen = (np.arange(npsf) - (npsf - 1) / 2.) / 2.

    def window(istar):
        lx, ux, ly, uy = _box_bounds(xc[istar], yc[istar], nhalf)
        if lx < 0 or ly < 0 or ux >= ncol or uy >= nrow:
            print('GETPSF: Star ', str(istar), ' too near edge of frame.')
            return None
        return np.asarray(image[int(ly):int(uy) + 1, int(lx):int(ux) + 1], dtype=np.float64), lx, ly

    def table_from(resid, x, y):
        gx, gy = np.meshgrid(psfgen + x, psfgen + y)
        return rinter.rinter(resid, gx, gy, deriv=False, device=device).reshape(npsf, npsf)

    # ---- first usable star: Gaussian core fit + first residual table (getpsf.py:158-306)
    psf = None
    k = -1
    while k + 1 < numpsf:
        k += 1
        istar = idpsf[k]
        w = window(istar)
        if w is None:
            continue
        f, lx, ly = w
        f = f - sky[istar]
        x, y = xc[istar] - lx, yc[istar] - ly
        if verbose:
            print('STAR  X  Y  MAG  SKY')
            print(istar, xc[istar], yc[istar], mag[istar], sky[istar])
            print('')
        fit = _fit_core_gaussian(f, x, y, device)
        if fit is None:
            print('No convergence after 100 iterations for star ' + str(istar))
            gauss = np.array([-1, -1, -1, -1, -1])
            if k >= numpsf - 1:
                return (gauss, -1, -1)
            continue
        h, dxcen, dycen, sigx, sigy = fit
        model, _ = daoerf.daoerf(xgen, ygen, np.array([h, x + dxcen, y + dycen, sigx, sigy]), device=device)
        psfmag = mag[istar]
        psf = table_from(f - model.reshape(nbox, nbox), x, y)
        gauss = np.array([h, dxcen, dycen, sigx, sigy])
        goodstar = [k]
        break
    if psf is None:
        if k >= 0 and numpsf:
            print('ERROR - No valid PSF stars were supplied')
        return (np.array([-1, -1, -1, -1, -1]), -1, -1)

    # ---- every further star is fitted with the table so far, then folded into it (getpsf.py:308-386)
    for k in range(k + 1, numpsf):
        istar = idpsf[k]
        with np.errstate(all='ignore'):
            scale = 10. ** (-0.4 * (mag[istar] - psfmag))
        w = window(istar)
        if w is None:
            continue
        f, lx, ly = w
        print('STAR  X  Y  MAG  SKY')
        print(istar, xc[istar], yc[istar], mag[istar], sky[istar])
        print('')
        x, y = xc[istar] - lx, yc[istar] - ly
        pk = pkfit.pkfit_class(f, gauss, psf, ronois, phpadu, device=device)
        errmag, chi, sharp, niter, scale, x, y = pk.pkfit(scale, x, y, sky[istar], fitrad, debug=debug, xyout=True)
        if niter == 25 or scale == 1000000.0:
            print('GETPSF: No convergence after 25 iterations or invalid scale for star', istar)
            continue
        model, _ = daoerf.daoerf(xgen, ygen, np.array([gauss[0], x + dxcen, y + dycen, sigx, sigy]), device=device)
        resid = f - scale * model.reshape(nbox, nbox) - sky[istar]
        psf = psf + _patch_nans(table_from(resid, x, y))
        psfmag = -2.5 * np.log10((1. + scale) * 10 ** (-0.4 * psfmag))
        gauss[0] = gauss[0] * (1. + scale)
        goodstar.append(k)

    if len(goodstar) == 1:
        print('Warning: Only one valid PSF star')
    if psfname:
        write_psf(psfname, gauss, psf, psfmag, psfrad, fitrad, phpadu, ronois, zeropoint, goodstar[-1], len(goodstar))
    return (gauss, psf, psfmag)


def write_psf(psfname, gauss, psf, psfmag, psfrad, fitrad, phpadu, ronois, zeropoint, psfid, nstars):
    """The PSF FITS file of getpsf.py:400-427 (same keywords, comments and data)."""
    cards = [('PHPADU', float(phpadu), 'Photons per Analog Digital Unit'), ('RONOIS', float(ronois), 'Readout Noise'),
             ('PSFRAD', float(psfrad), 'Radius where PSF is defined (pixels)'), ('FITRAD', float(fitrad), 'Fitting Radius'),
             ('PSFMAG', float(psfmag), 'PSF Magnitude'), ('GAUSS1', float(gauss[0]), 'Gaussian Scale Factor'),
             ('GAUSS2', float(gauss[1]), 'Gaussian X Position'), ('GAUSS3', float(gauss[2]), 'Gaussian Y Position'),
             ('GAUSS4', float(gauss[3]), 'Gaussian Sigma: X Direction'), ('GAUSS5', float(gauss[4]), 'Gaussian Sigma: Y Direction'),
             ('PSFZPT', float(zeropoint), 'zeropoint used to scale psf mag'), ('PSFID', int(psfid)), ('NSTARS', int(nstars))]
    _fits.write_image(psfname, np.asarray(psf, dtype=np.float64), cards)
