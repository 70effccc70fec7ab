"""CPU restatement of PythonPhot/find.py (DAOFIND) -- TEST INFRASTRUCTURE ONLY.

Nothing in the product package imports this file; tests/test_find.py uses it as the checker for
csrc/find_kernel.cu on fields larger than the committed golden vectors.  It is pinned to the
unmodified reference by tests/golden/golden_find.npz (tests/test_find.py, CPU leg).

Restated with numpy, vectorised over candidates instead of the reference's per-candidate Python loop;
the convolution is the same third-party call the reference makes (scipy.ndimage.convolve, find.py:5,188).
"""
import numpy as np
from scipy.ndimage import convolve


def find_constants(fwhm, maxbox=13):
    """Everything find.py:106-183 derives from fwhm alone."""
    radius = max(0.637 * fwhm, 2.001)  # find.py:106-107
    radsq = radius ** 2
    nhalf = min(int(radius), int((maxbox - 1) / 2.))
    nbox = 2 * nhalf + 1
    sigsq = (fwhm / 2.35482) ** 2
    off = np.arange(nbox) - nhalf
    row2 = off ** 2
    dist2 = (row2[:, None] + row2[None, :]).astype(np.float64)  # find.py:123-126
    mask = (dist2 <= radsq).astype(np.int8)
    pixels = int(mask.sum())
    g = np.exp(-0.5 * dist2 / sigsq)
    wt = (nhalf - np.abs(off) + 1).astype(np.float64)  # find.py:150-152
    xwt = np.tile(wt, (nbox, 1))
    ywt = xwt.T
    K = dict(nhalf=nhalf, nbox=nbox, sigsq=sigsq, wt=wt)
    K["sgx"] = np.sum(g * xwt, 1)
    K["p"] = np.sum(wt)
    K["sgy"] = np.sum(g * ywt, 0)
    K["sumgx"] = np.sum(wt * K["sgy"])
    K["sumgy"] = np.sum(wt * K["sgx"])
    K["sumgsqy"] = np.sum(wt * K["sgy"] * K["sgy"])
    K["sumgsqx"] = np.sum(wt * K["sgx"] * K["sgx"])
    vec = (nhalf - np.arange(nbox)).astype(np.float64)
    K["dgdx"] = K["sgy"] * vec
    K["dgdy"] = K["sgx"] * vec
    K["sdgdxs"] = np.sum(wt * K["dgdx"] ** 2)
    K["sdgdx"] = np.sum(wt * K["dgdx"])
    K["sdgdys"] = np.sum(wt * K["dgdy"] ** 2)
    K["sdgdy"] = np.sum(wt * K["dgdy"])
    K["sgdgdx"] = np.sum(wt * K["sgy"] * K["dgdx"])
    K["sgdgdy"] = np.sum(wt * K["sgx"] * K["dgdy"])
    c = g * mask  # find.py:174-179
    sumc = np.sum(c)
    sumcsq = np.sum(c ** 2) - sumc ** 2 / pixels
    sumc = sumc / pixels
    c[mask == 1] = (c[mask == 1] - sumc) / sumcsq
    K["c"] = c
    c1 = np.exp(-.5 * row2 / sigsq)  # find.py:180-183
    sumc1 = np.sum(c1) / nbox
    sumc1sq = np.sum(c1 ** 2) - sumc1
    K["c1"] = (c1 - sumc1) / sumc1sq
    ring = mask.copy()
    ring[nhalf, nhalf] = 0  # find.py:199-200: the centre pixel is excluded from here on
    K["ring"] = ring
    K["pixels"] = pixels - 1
    return K


def find(image, hmin, fwhm, roundlim=(-1.0, 1.0), sharplim=(0.2, 1.0), return_counts=False):
    """find.py:89-394.  Returns (x, y, flux, sharp, roundness) or None (plus a counts dict on request)."""
    image = np.asarray(image).astype(np.float64)
    n_y, n_x = image.shape
    K = find_constants(fwhm)
    nhalf, nbox, pixels = K["nhalf"], K["nbox"], K["pixels"]
    h = convolve(image, K["c"])  # find.py:188
    minh = np.min(h)
    h[:, 0:nhalf] = minh; h[:, n_x - nhalf:n_x] = minh  # find.py:191-192
    h[0:nhalf, :] = minh; h[n_y - nhalf:n_y - 1, :] = minh  # sic: the last row keeps its convolved values
    ry, rx = np.nonzero(K["ring"])
    yy, xx = ry - nhalf, rx - nhalf
    iy, ix = np.nonzero(h >= hmin)  # find.py:211
    for k in range(pixels):  # find.py:220-233: keep pixels that are >= every ring neighbour, all of them in the frame
        hy, hx = iy + yy[k], ix + xx[k]
        ok = (hy < n_y) & (hx < n_x) & (hy >= 0) & (hx >= 0)
        iy, ix, hy, hx = iy[ok], ix[ok], hy[ok], hx[ok]
        keep = h[iy, ix] >= h[hy, hx]
        iy, ix = iy[keep], ix[keep]
    ngood = iy.size
    counts = dict(ngood=ngood, badsharp=0, badround=0, badcntrd=0)
    if ngood == 0:
        return (None, counts) if return_counts else None
    # candidate boxes [ngood, nbox, nbox] (every candidate is >= nhalf pixels from the frame edges here)
    o = np.arange(nbox) - nhalf
    temp = image[(iy[:, None] + o[None, :])[:, :, None], (ix[:, None] + o[None, :])[:, None, :]]
    d = h[iy, ix]
    with np.errstate(all="ignore"):
        sharp1 = (temp[:, nhalf, nhalf] - np.sum(K["ring"][None] * temp, axis=(1, 2)) / pixels) / d  # find.py:265
        alive = ~((sharp1 < sharplim[0]) | (sharp1 > sharplim[1]))
        counts["badsharp"] = int((~alive).sum())
        dx = np.sum(np.sum(temp, axis=1) * K["c1"][None], axis=1)  # find.py:272-273
        dy = np.sum(np.sum(temp, axis=2) * K["c1"][None], axis=1)
        bad = alive & ((dx <= 0) | (dy <= 0))
        around = 2 * (dx - dy) / (dx + dy)
        bad |= alive & ~bad & ((around < roundlim[0]) | (around > roundlim[1]))
        counts["badround"] = int(bad.sum())
        alive &= ~bad
        wt, p, sigsq = K["wt"], K["p"], K["sigsq"]
        # x centroid, find.py:293-313
        sd = np.sum(temp * wt[None, :, None], axis=1)
        sumgd = np.sum(wt * K["sgy"] * sd, axis=1)
        sumd = np.sum(wt * sd, axis=1)
        sddgdx = np.sum(wt * sd * K["dgdx"], axis=1)
        hx = (sumgd - K["sumgx"] * sumd / p) / (K["sumgsqy"] - K["sumgx"] ** 2 / p)
        bad = alive & (hx <= 0)
        skylvl = (sumd - hx * K["sumgx"]) / p
        ddx = (K["sgdgdx"] - (sddgdx - K["sdgdx"] * (hx * K["sumgx"] + skylvl * p))) / (hx * K["sdgdxs"] / sigsq)
        bad |= alive & ~bad & (np.abs(ddx) >= nhalf)
        alive &= ~bad
        nbadc = int(bad.sum())
        # y centroid, find.py:317-338
        sd = np.sum(temp * wt[None, None, :], axis=2)
        sumgd = np.sum(wt * K["sgx"] * sd, axis=1)
        sumd = np.sum(wt * sd, axis=1)
        sddgdy = np.sum(wt * sd * K["dgdy"], axis=1)
        hyv = (sumgd - K["sumgy"] * sumd / p) / (K["sumgsqx"] - K["sumgy"] ** 2 / p)
        bad = alive & (hyv <= 0)
        skylvl = (sumd - hyv * K["sumgy"]) / p
        ddy = (K["sgdgdy"] - (sddgdy - K["sdgdy"] * (hyv * K["sumgy"] + skylvl * p))) / (hyv * K["sdgdys"] / sigsq)
        bad |= alive & ~bad & (np.abs(ddy) >= nhalf)
        alive &= ~bad
        counts["badcntrd"] = nbadc + int(bad.sum())
    if not alive.any():
        return (None, counts) if return_counts else None
    res = (ix[alive] + ddx[alive], iy[alive] + ddy[alive], d[alive], sharp1[alive], around[alive])
    return (res, counts) if return_counts else res
