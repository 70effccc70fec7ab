"""Golden vectors for getpsf.getpsf and find.find from the UNMODIFIED reference (authoring container only):

    python tests/golden/gen_golden_psf_find.py

Same rules as gen_golden.py: the reference is imported in-process through oracle/refloader.py, called with
fp64 arrays, and its inputs are stored beside its outputs in ``golden_getpsf.npz`` / ``golden_find.npz``.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import refloader  # noqa: E402
from pythonphot_b200 import synth  # noqa: E402

GETPSF_CASES = [  # name, psfrad, fitrad, which stars (indices into the bright-star list)
    ("r8_f3", 8, 3, None),
    ("r6_f2p5", 6, 2.5, None),
    ("single_star", 8, 3, [2]),
    ("edge_star_first", 8, 3, "edge_first"),
    ("fitrad_too_big", 4, 5, None),
]
FIND_CASES = [  # name, hmin, fwhm, roundlim, sharplim
    ("fwhm4", 60.0, 4.1, [-1.0, 1.0], [0.2, 1.0]),
    ("fwhm2p5_loose", 40.0, 2.5, [-2.0, 2.0], [0.1, 1.2]),
    ("fwhm9_bigbox", 80.0, 9.0, [-1.0, 1.0], [0.2, 1.0]),
    ("none_found", 1e9, 4.1, [-1.0, 1.0], [0.2, 1.0]),
]


def psf_field():
    """A sparse field of bright stars: several usable PSF stars, one too close to the edge, a few NaN pixels
    inside one star's window (exercises the NaN patching of getpsf.py:357-373)."""
    f = synth.make_field(260, 280, 26, seed=21, dtype=np.float64, margin=20.0, edge_frac=0.0, nan_frac=0.0, sat_frac=0.0,
                         scale_range=(30.0, 120.0), with_noise=False)
    img = f["image"].copy()
    x, y = f["x"].copy(), f["y"].copy()
    # star 4 is moved next to the frame edge in the LIST (its window leaves the frame -> skipped)
    x[4], y[4] = 5.3, 100.2
    # NaNs in the outskirts of star 6's window
    ix, iy = int(x[6]), int(y[6])
    img[iy + 7, ix - 6] = np.nan
    img[iy - 8, ix + 5] = np.nan
    img[iy - 8, ix + 6] = np.nan
    return img, x, y, f


def gen_getpsf():
    getpsf = refloader.load("getpsf")
    aper = refloader.load("aper")
    img, x, y, f = psf_field()
    with refloader.quiet(), np.errstate(all="ignore"):
        r = aper.aper(img, x, y, phpadu=f["phpadu"], apr=np.array([5.0]), zeropoint=25.0, skyrad=[15, 22], badpix=[-12000, 60000],
                      exact=False, verbose=False)
    mag, sky = np.asarray(r[0], float).reshape(-1), np.asarray(r[4], float).reshape(-1)
    mag = np.where(np.isfinite(mag), mag, 99.0)
    out = dict(image=img, x=x, y=y, mag=mag, sky=sky, ronois=np.float64(f["ronois"]), phpadu=np.float64(f["phpadu"]))
    order = np.argsort(mag)  # brightest first, as a user would pick PSF stars
    for name, psfrad, fitrad, sel in GETPSF_CASES:
        o = [int(v) for v in order if v not in (4, 6)]
        if sel is None:  # bright stars, with the edge star (4) and the NaN-window star (6) in the middle of the list
            idpsf = np.array(o[:3] + [4] + o[3:8] + [6] + o[8:10])
        elif sel == "edge_first":
            idpsf = np.array([4, o[0], 6, o[1]])
        else:
            idpsf = order[np.array(sel)]
        with refloader.quiet() as log, np.errstate(all="ignore"):
            g, p, m = getpsf.getpsf(img, x, y, mag, sky, f["ronois"], f["phpadu"], idpsf, psfrad, fitrad, "unused.fits", verbose=False)
        out[name + "__idpsf"] = np.asarray(idpsf)
        out[name + "__gauss"] = np.asarray(g, dtype=np.float64)
        out[name + "__psf"] = np.asarray(p, dtype=np.float64)
        out[name + "__psfmag"] = np.float64(m)
        out[name + "__nskipped"] = np.int64(log.getvalue().count("too near edge") + log.getvalue().count("No convergence"))
    return out


def gen_find():
    find = refloader.load("find")
    f = synth.make_field(300, 340, 260, seed=31, dtype=np.float64, margin=4.0, edge_frac=0.05, nan_frac=0.0, sat_frac=1e-4,
                         scale_range=(1.0, 120.0), with_noise=False)
    img = f["image"]
    out = dict(image=img)
    for name, hmin, fwhm, roundlim, sharplim in FIND_CASES:
        with refloader.quiet(), np.errstate(all="ignore"):
            r = find.find(img, hmin, fwhm, roundlim, sharplim, verbose=False)
        if r is None:
            r = [np.zeros(0)] * 5
            out[name + "__none"] = np.int64(1)
        for nm, a in zip(("x", "y", "flux", "sharp", "round"), r):
            out["%s__%s" % (name, nm)] = np.asarray(a, dtype=np.float64)
    # NaN pixels: h is NaN around them, min(h) is NaN, so the border assignment (find.py:192-194) writes NaN
    rng = np.random.default_rng(32)
    nan_idx = rng.integers(0, img.size, 12)
    img2 = img.copy()
    img2.reshape(-1)[nan_idx] = np.nan
    out["nan_idx"] = nan_idx
    with refloader.quiet(), np.errstate(all="ignore"):
        r = find.find(img2, 60.0, 4.1, [-1.0, 1.0], [0.2, 1.0], verbose=False)
    for nm, a in zip(("x", "y", "flux", "sharp", "round"), r):
        out["with_nans__%s" % nm] = np.asarray(a, dtype=np.float64)
    return out


CNTRD_CASES = [  # name, fwhm, extendbox, keepcenter
    ("fwhm4", 4.1, False, False),
    ("fwhm4_extend3", 4.1, 3, False),
    ("fwhm4_extend_xy", 4.1, (1, 4), False),
    ("fwhm7_keepcenter", 7.0, False, True),
    ("fwhm2", 2.0, False, False),
]


def gen_cntrd():
    """cntrd.cntrd on approximate star positions: jittered truth, frame-edge positions, a rounded (integer-valued,
    tie-prone) copy of the frame, NaN pixels near two stars."""
    cntrd = refloader.load("cntrd")
    f = synth.make_field(220, 260, 180, seed=41, dtype=np.float64, margin=3.0, edge_frac=0.1, nan_frac=0.0, sat_frac=2e-4,
                         scale_range=(3.0, 150.0), with_noise=False)
    rng = np.random.default_rng(42)
    x = f["x"] + rng.normal(0, 0.8, f["x"].size)
    y = f["y"] + rng.normal(0, 0.8, f["y"].size)
    img = f["image"].copy()
    for k in (3, 17):
        img[int(round(f["y"][k])) + 1, int(round(f["x"][k])) - 1] = np.nan
    out = dict(image=img, x=x, y=y)
    for tag, frame in (("", img), ("rounded__", np.round(img / 50.0) * 50.0)):
        for name, fwhm, ext, keep in CNTRD_CASES:
            with refloader.quiet(), np.errstate(all="ignore"):
                xc, yc = cntrd.cntrd(frame, x.copy(), y.copy(), fwhm, verbose=False, extendbox=ext, keepcenter=keep)
            out[tag + name + "__xcen"] = np.asarray(xc, dtype=np.float64)
            out[tag + name + "__ycen"] = np.asarray(yc, dtype=np.float64)
    k = int(np.flatnonzero(out["fwhm4__xcen"] > 0)[0])
    with refloader.quiet():
        xs, ys = cntrd.cntrd(img, float(x[k]), float(y[k]), 4.1, verbose=False)
    out["scalar__k"] = np.int64(k)
    out["scalar__xy"] = np.array([xs, ys], dtype=np.float64)
    return out



SIGCLIP_CASES = [  # name, aper keywords; inputs are golden_aper.npz's image and star list
    ("exact_4050", dict(apr=np.array([3.0, 5.0, 8.0]), skyrad=[40, 50], badpix=[-12000, 60000], exact=True)),
    ("approx_1220_nobadpix", dict(apr=np.array([2.5, 6.0]), skyrad=[12, 20], badpix=[0, 0], exact=False)),
]


def gen_aper_sigclip():
    """aper.aper(skyalgorithm='sigmaclipping') (aper.py:275-282) on the frame of golden_aper.npz (fp64 copy)."""
    aper = refloader.load("aper")
    g = np.load(os.path.join(HERE, "golden_aper.npz"))
    img64, x, y = g["image"].astype(np.float64), g["x"], g["y"]
    out = {}
    for name, kw in SIGCLIP_CASES:
        with refloader.quiet(), np.errstate(all="ignore"):
            r = aper.aper(img64, x, y, phpadu=float(g["phpadu"]), zeropoint=25.0, verbose=False, skyalgorithm="sigmaclipping", **kw)
        for nm, a in zip(("mag", "magerr", "flux", "fluxerr", "sky", "skyerr", "badflag"), r[:7]):
            a = np.asarray(a, dtype=np.float64)
            out["%s__%s" % (name, nm)] = a.reshape(-1) if nm in ("sky", "skyerr") else a.reshape(x.size, -1)
    return out


def main(which=("getpsf", "find", "cntrd", "sigclip")):
    if "getpsf" in which:
        np.savez_compressed(os.path.join(HERE, "golden_getpsf.npz"), **gen_getpsf())
    if "find" in which:
        np.savez_compressed(os.path.join(HERE, "golden_find.npz"), **gen_find())
    if "cntrd" in which:
        np.savez_compressed(os.path.join(HERE, "golden_cntrd.npz"), **gen_cntrd())
    if "sigclip" in which:
        np.savez_compressed(os.path.join(HERE, "golden_aper_sigclip.npz"), **gen_aper_sigclip())
    for fn in sorted(os.listdir(HERE)):
        if fn.endswith(".npz"):
            print(fn, os.path.getsize(os.path.join(HERE, fn)))


if __name__ == "__main__":
    main(sys.argv[1:] or ("getpsf", "find", "cntrd", "sigclip"))
