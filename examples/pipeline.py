"""The PythonPhot notebook pipeline (examples/PythonPhot.ipynb of the reference: find -> cntrd -> aper -> getpsf ->
pkfit) on a synthetic frame, every step through pythonphot_b200's drop-in modules (CUDA).

    python examples/pipeline.py [size] [nstars]
"""
import contextlib
import io
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from pythonphot_b200 import aper, cntrd, find, getpsf, pkfit_noise, synth  # noqa: E402


def run(size=2048, nstars=6000, seed=1001, verbose=True):
    f = synth.make_field(size, size, nstars, seed=seed, dtype=np.float32, margin=30.0, edge_frac=0.0, nan_frac=0.0, sat_frac=0.0)
    image, t = f["image"], {}
    say = print if verbose else (lambda *a, **k: None)
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        xf, yf, flux, sharp, rnd = find.find(image, 60.0, 4.1, verbose=False)
    t["find"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    xc, yc = cntrd.cntrd(image, xf, yf, 4.1, verbose=False)
    ok = xc > 0
    xc, yc = xc[ok], yc[ok]
    t["cntrd"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    mag, magerr, fl, flerr, sky, skyerr, bad, _ = aper.aper(image, xc, yc, phpadu=f["phpadu"], apr=5, zeropoint=25, skyrad=[40, 50],
                                                            badpix=[-12000, 60000], exact=True, verbose=False)
    t["aper"] = time.perf_counter() - t0
    # PSF stars: unflagged, with no other detection within 15 px (a blend as first star would set the Gaussian widths)
    from scipy.spatial import cKDTree
    dnn = cKDTree(np.column_stack([xc, yc])).query(np.column_stack([xc, yc]), k=2)[0][:, 1]
    good = np.flatnonzero((bad.reshape(xc.size, -1)[:, 0] == 0) & np.isfinite(mag.reshape(xc.size, -1)[:, 0]) & (dnn > 15.0))
    order = good[np.argsort(mag.reshape(xc.size, -1)[good, 0])]
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        gauss, psf, psfmag = getpsf.getpsf(image.astype(np.float64), xc, yc, mag.reshape(xc.size, -1)[:, 0], sky, f["ronois"], f["phpadu"],
                                           order[:15], 8, 3, None, verbose=False)
    t["getpsf"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    pk = pkfit_noise.pkfit_class(image, gauss, psf, f["ronois"], f["phpadu"], f["noise"], f["mask"])
    errmag, chi, shp, niter, scale, xfit, yfit = pk.pkfit_noise(np.ones(xc.size), xc, yc, sky, 5.0, xyout=True)
    t["pkfit_noise"] = time.perf_counter() - t0
    say("%d x %d frame, %d stars planted: find %d, centroided %d, aper ok %d, PSF sigma %.3f/%.3f, fits converged %d" % (
        size, size, nstars, xf.size, xc.size, good.size, gauss[3], gauss[4], int(((niter > 0) & (niter < 25)).sum())))
    say("seconds: " + ", ".join("%s %.3f" % kv for kv in t.items()))
    return dict(found=(xf, yf), xy=(xc, yc), mag=mag, sky=sky, gauss=gauss, psf=psf, psfmag=psfmag, scale=scale, niter=niter,
                xyfit=(xfit, yfit), field=f)


if __name__ == "__main__":
    run(int(sys.argv[1]) if len(sys.argv) > 1 else 2048, int(sys.argv[2]) if len(sys.argv) > 2 else 6000)
