"""CPU restatement of aper.py's sigma-clipped sky (skyalgorithm='sigmaclipping', aper.py:240-282) -- TEST INFRASTRUCTURE ONLY.

Pinned to the unmodified reference by tests/golden/golden_aper_sigclip.npz; calls the same third-party routine the
reference calls (scipy.stats.sigmaclip)."""
import numpy as np
from scipy.stats import sigmaclip


def sky(image, xc, yc, skyrad, badpix, minsky=20):
    """Per star: (skymod, skysig, nsky); stars off the frame or with too few sky pixels give (0, 0, nsky)."""
    image = np.asarray(image, dtype=np.float64)
    nrow, ncol = image.shape
    xc, yc = np.atleast_1d(np.asarray(xc, float)), np.atleast_1d(np.asarray(yc, float))
    rin2, rout2 = float(skyrad[0]) ** 2, float(skyrad[1]) ** 2
    out = np.zeros((xc.size, 3))
    for i, (x, y) in enumerate(zip(xc, yc)):
        if x < 0.5 or x > ncol - 1.5 or y < 0.5 or y > nrow - 1.5:  # aper.py:209-210
            continue
        lx, ux = max(int(x - skyrad[1]), 0), min(int(x + skyrad[1]), ncol - 1)  # aper.py:188-199
        ly, uy = max(int(y - skyrad[1]), 0), min(int(y + skyrad[1]), nrow - 1)
        stamp = image[ly:uy + 1, lx:ux + 1]
        rsq = (np.arange(ux - lx + 1)[None, :] - (x - lx)) ** 2 + (np.arange(uy - ly + 1)[:, None] - (y - ly)) ** 2
        sel = (rsq >= rin2) & (rsq <= rout2)
        buf = stamp[sel]
        if badpix[0] != badpix[1]:  # aper.py:260-266 (strict; NaNs survive)
            buf = buf[~((buf < badpix[0]) | (buf > badpix[1]))]
        out[i, 2] = buf.size
        if buf.size < minsky:
            continue
        with np.errstate(all="ignore"):
            c = sigmaclip(buf, low=4.0, high=4.0)[0]
            out[i, 0], out[i, 1] = np.mean(c), np.std(c)
    sig = np.where(out[:, 1] > 999.99, 999.0, out[:, 1])  # aper.py:298, applied to the stored skyerr only
    return out[:, 0], sig, out[:, 2].astype(int)
